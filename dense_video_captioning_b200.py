"""Import alias for the package directory `dense-video-captioning_b200/`.

The repository layout names the package with a hyphen, which Python cannot import directly; this
one-file shim loads that directory as the package `dense_video_captioning_b200` (sub-packages and
relative imports work as usual):

    import dense_video_captioning_b200 as dvc
    from dense_video_captioning_b200.ops.modules import MSDeformAttn
"""
import importlib.util
import os
import sys

_dir = os.path.join(os.path.dirname(os.path.abspath(__file__)), "dense-video-captioning_b200")
_spec = importlib.util.spec_from_file_location(__name__, os.path.join(_dir, "__init__.py"),
                                               submodule_search_locations=[_dir])
_mod = importlib.util.module_from_spec(_spec)
sys.modules[__name__] = _mod
_spec.loader.exec_module(_mod)
