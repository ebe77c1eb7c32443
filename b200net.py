"""Import shim: the package directory carries the (hyphenated) project name required by the repository layout;
`import b200net` loads it under a valid Python module name."""
import importlib.util
import os
import sys

_dir = os.path.join(os.path.dirname(os.path.abspath(__file__)),
                    "performance-comparison-of-tensorflow-pytorch-and-their-distributed-counterparts_b200")
_spec = importlib.util.spec_from_file_location("b200net", os.path.join(_dir, "__init__.py"),
                                               submodule_search_locations=[_dir])
_mod = importlib.util.module_from_spec(_spec)
sys.modules["b200net"] = _mod
_spec.loader.exec_module(_mod)
