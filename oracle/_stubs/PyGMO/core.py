class champion:
    def __init__(self):
        self.x = ()
        self.f = ()
        self.c = ()
