"""montecarlo_b200 — B200-native (sm_100a, fp64 CUDA) trial-move energy path of MatthijsLasure/MonteCarlo.

  boxio   the Fortran driver's input formats (box / DMSO / param files), box replication
  engine  LJEngine: numpy <-> C ABI (include/mcb200.h) of libmcb200.so; all arithmetic runs on the GPU
"""
from . import boxio  # noqa: F401


def __getattr__(name):  # engine needs the built CUDA library; keep `import montecarlo_b200` cheap
    if name in ("LJEngine", "McbError", "measure_fp64_peak", "engine"):
        import importlib
        engine = importlib.import_module(".engine", __name__)   # (not `from . import`: that asks __getattr__ again)
        return engine if name == "engine" else getattr(engine, name)
    raise AttributeError(name)
