"""Type aliases used across the package (same names as the toolbox's
source/core/utilities/type_aliases.py)."""
from typing import Tuple

import numpy as np

Vector = np.ndarray
Float = np.float64
Tuple2V = Tuple[Vector, Vector]
Tuple3V = Tuple[Vector, Vector, Vector]
Tuple4V = Tuple[Vector, Vector, Vector, Vector]
