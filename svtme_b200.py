"""Import shim: the package directory is named `svt-av1-psy_b200` (not a Python identifier), so
it is loaded by path and registered as `svt_av1_psy_b200`. `import svtme_b200 as me` gives the
package; `from svt_av1_psy_b200 import abi` works after this module has been imported."""
import importlib.util
import os
import sys

_NAME = "svt_av1_psy_b200"
_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "svt-av1-psy_b200")

if _NAME not in sys.modules:
    _spec = importlib.util.spec_from_file_location(
        _NAME, os.path.join(_DIR, "__init__.py"), submodule_search_locations=[_DIR]
    )
    _mod = importlib.util.module_from_spec(_spec)
    sys.modules[_NAME] = _mod
    _spec.loader.exec_module(_mod)

package = sys.modules[_NAME]
globals().update({k: v for k, v in vars(package).items() if not k.startswith("__")})
