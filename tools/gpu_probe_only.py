"""DEVELOPMENT TOOL: runs only the FP64 peak probe (for an ncu capture of the roofline's denominator)."""
import os, sys
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
from lassim_b200 import Engine
eng = Engine(0)
print("fp64 peak probe: %.2f TFLOP/s" % eng.fp64_peak_tflops(400000))
