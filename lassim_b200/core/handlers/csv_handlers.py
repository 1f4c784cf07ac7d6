"""CSV solutions handler (toolbox: source/core/handlers/csv_handlers.py)."""
import os

from ..solutions_handler import SolutionsHandler
from .serializers.csv_serializer import CSVSerializer


def default_filename_creator(num_solutions, num_variables):
    return "top{}solutions_{}variables_{}.csv".format(num_solutions, num_variables, os.getpid())


class SimpleCSVSolutionsHandler(SolutionsHandler):
    def __init__(self, output_dir, num_solutions, headers,
                 filename_creator=default_filename_creator):
        self.csv_serializer = CSVSerializer.new_instance(
            output_dir, num_solutions, headers, filename_creator)

    def handle_solutions(self, solutions, filename=None):
        self.csv_serializer.serialize_solutions(solutions, filename)
