"""Python face of the C++ host runtime (csrc/core, bound in csrc/bindings/core_bindings.cpp).

    from minips_b200 import core
    e = core.Engine(core.Node(0, "localhost", 14560), [core.Node(0, "localhost", 14560)])
    e.start_everything(); tid = e.create_table([(0, 100)], core.ModelType.BSP, core.StorageType.Vector)

The extension is built in-tree by build.py (minips_b200/lib/_minips_core*.so).
"""
import importlib.util
import sysconfig
from pathlib import Path

_path = Path(__file__).resolve().parent / "lib" / f"_minips_core{sysconfig.get_config_var('EXT_SUFFIX')}"
if not _path.exists():
    spec = importlib.util.spec_from_file_location("minips_build", _path.parents[2] / "build.py")
    _b = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(_b)
    _b.build_core(verbose=False)
if not _path.exists():
    raise ImportError(f"{_path} not built; run `python build.py --only core`")
_spec = importlib.util.spec_from_file_location("_minips_core", _path)
_mod = importlib.util.module_from_spec(_spec)
_spec.loader.exec_module(_mod)
globals().update({k: getattr(_mod, k) for k in dir(_mod) if not k.startswith("__")})
