"""Migration topologies: who receives an island's best individual."""


class unconnected:
    def neighbours(self, i, n):
        return []


class ring:
    """Bidirectional ring as in PyGMO's topology.ring: island i sends to its
    two neighbours."""

    def neighbours(self, i, n):
        if n < 2:
            return []
        if n == 2:
            return [(i + 1) % n]
        return [(i - 1) % n, (i + 1) % n]


class one_way_ring:
    def neighbours(self, i, n):
        return [(i + 1) % n] if n > 1 else []


class fully_connected:
    def neighbours(self, i, n):
        return [j for j in range(n) if j != i]
