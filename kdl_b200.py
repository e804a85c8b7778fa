"""Import shim: the product package lives in the directory ``kubernetes-deep-learning_b200/`` (the name the
project layout prescribes), which is not a legal Python identifier.  ``import kdl_b200`` loads that
directory as the package ``kdl_b200``."""
import importlib.util
import os
import sys

_dir = os.path.join(os.path.dirname(os.path.abspath(__file__)), "kubernetes-deep-learning_b200")
_spec = importlib.util.spec_from_file_location(
    "kdl_b200", os.path.join(_dir, "__init__.py"), submodule_search_locations=[_dir])
_mod = importlib.util.module_from_spec(_spec)
sys.modules["kdl_b200"] = _mod
_spec.loader.exec_module(_mod)
