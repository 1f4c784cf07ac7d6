class champion:
    """Best individual of a population: decision vector x, fitness f (a
    1-tuple), constraints c (unused)."""

    def __init__(self, x=(), f=(), c=()):
        self.x = tuple(x)
        self.f = tuple(f)
        self.c = tuple(c)

    def __repr__(self):
        return "champion(f={}, dim={})".format(self.f, len(self.x))
