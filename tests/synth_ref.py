"""numpy restatement of the device generators; lives in the package (s-blas_b200/synth_np.py)
because bench.py's reference arm builds its inputs with it without touching CUDA."""
import importlib.util
import os

_p = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "s-blas_b200", "synth_np.py")
_s = importlib.util.spec_from_file_location("sblas_synth_np", _p)
_m = importlib.util.module_from_spec(_s)
_s.loader.exec_module(_m)
globals().update({k: v for k, v in vars(_m).items() if not k.startswith("__")})
