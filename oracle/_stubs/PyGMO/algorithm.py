class _Algo:
    def __init__(self, **kw):
        self.kw = kw


de = jde = mde_pbx = de_1220 = pso = pso_gen = sga_gray = sa_corana = _Algo
bee_colony = cmaes = _Algo
