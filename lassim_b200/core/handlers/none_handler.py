from ..solutions_handler import SolutionsHandler


class NoneSolutionsHandler(SolutionsHandler):
    def handle_solutions(self, list_of_solutions):
        pass
