from ..solutions_handler import SolutionsHandler


class CompositeSolutionsHandler(SolutionsHandler):
    """Forwards every list of solutions to each handler in turn (toolbox:
    source/core/handlers/composite_handler.py)."""

    def __init__(self, handlers_list):
        self._handlers = handlers_list

    def handle_solutions(self, list_of_solutions):
        for handler in self._handlers:
            handler.handle_solutions(list_of_solutions)
