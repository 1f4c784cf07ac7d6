"""Network -> core system (sorted transcription factors, reversed reaction
map).  Behaviour follows the toolbox's source/core/core_system.py:13-133."""
from sortedcontainers import SortedDict, SortedSet


class CoreSystem:
    """`network` maps each transcription factor to the set of factors it acts
    on.  `reactions` is the reverse view: for each factor, who regulates it.
    With `correction`, a factor nobody regulates is connected to every factor,
    itself included (core_system.py:55-60)."""

    def __init__(self, network, correction=True):
        self._network = network
        names = set(network.keys())
        for targets in network.values():
            names.update(targets)
        self.tfacts = SortedSet(names)
        self.__reactions, self.__react_count = self._reverse(correction)

    def _reverse(self, correction):
        incoming = SortedDict()
        count = 0
        for source, targets in self._network.items():
            incoming.setdefault(source, SortedSet())
            for target in targets:
                incoming.setdefault(target, SortedSet()).add(source)
                count += 1
        if correction:
            for tfact in list(incoming.keys()):
                if len(incoming[tfact]) == 0:
                    incoming[tfact] = SortedSet(self.tfacts)
                    count += len(self.tfacts)
        return incoming, count

    def from_tfacts_to_ids(self):
        for idx, name in enumerate(self.tfacts):
            yield name, idx

    def from_ids_to_tfacts(self):
        for idx, name in enumerate(self.tfacts):
            yield idx, name

    def from_reactions_to_ids(self):
        ids = dict(self.from_tfacts_to_ids())
        for tfact, sources in self.reactions.items():
            yield ids[tfact], SortedSet(ids[s] for s in sources)

    def from_reactions_ids_to_str(self, reactions):
        names = dict(self.from_ids_to_tfacts())
        for key, value in reactions.items():
            try:
                yield names[key], SortedSet(names[v] for v in value)
            except (KeyError, TypeError):
                raise AttributeError(
                    "Value in key and/or value not present in core\nkey -> {}\n"
                    "value -> {}".format(key, value))

    @property
    def num_tfacts(self):
        return len(self.tfacts)

    @property
    def reactions(self):
        return self.__reactions

    @property
    def react_count(self):
        return self.__react_count

    def __str__(self):
        lines = ["== Core ==", "= List of transcription factors =",
                 ", ".join(self.tfacts), "= List of reactions ="]
        lines += [key + " --> " + ", ".join(value)
                  for key, value in self.reactions.items()]
        return "\n".join(lines)
