class unconnected:
    pass


class ring:
    pass
