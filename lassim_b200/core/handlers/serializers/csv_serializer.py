"""Solutions -> tab-separated text (toolbox:
source/core/handlers/serializers/csv_serializer.py).  Per solution: a header
row, one row per transcription factor `lambda vmax <regulation row>`, then
`Cost<TAB>value`."""
import logging
import os


class CSVSerializer:
    def __init__(self, directory, n_print, headers, filename_creator):
        self._directory = directory
        self._num = n_print
        self._headers = headers
        self._filename_creator = filename_creator
        self._separator = "\t"

    @classmethod
    def new_instance(cls, output_dir, n_print, headers, filename_creator):
        """Creates `output_dir` when missing; an empty header list is an error
        (csv_serializer.py:39-48)."""
        if not os.path.isdir(output_dir):
            os.makedirs(output_dir)
        if len(headers) == 0:
            raise RuntimeError("List of headers for solutions serialization is empty.")
        return CSVSerializer(output_dir, n_print, headers, filename_creator)

    def serialize_solutions(self, solutions, filename=None):
        """Write the best `n_print` solutions (fewer if fewer are given) to
        one file; its name comes from the filename creator unless given."""
        if filename is None:
            filename = self._filename_creator(self._num, solutions[0].number_of_variables)
        count = min(len(solutions), self._num)
        with open(os.path.join(self._directory, filename), "w") as out:
            for solution in solutions[:count] if count != float("inf") else solutions:
                matrix = solution.get_solution_matrix(headers=self._headers)
                out.write(matrix.to_csv(sep=self._separator, index=False))
                out.write("Cost{}{}\n".format(self._separator, solution.cost))
        logging.getLogger(__name__).info(
            "Created file {} in directory {}".format(filename, self._directory))
