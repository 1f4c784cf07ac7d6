class SolutionsHandler:
    """Receives the sorted solutions of every optimisation iteration
    (toolbox: source/core/solutions_handler.py)."""

    def handle_solutions(self, list_of_solutions):
        raise NotImplementedError(self.handle_solutions.__name__)
