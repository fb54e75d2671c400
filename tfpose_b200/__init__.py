"""Import alias: ``import tfpose_b200`` loads the package that lives in the
directory ``tf-pose-estimation_b200/`` (a hyphen is not importable by name)."""
import os as _os

_real = _os.path.normpath(_os.path.join(_os.path.dirname(_os.path.abspath(__file__)), "..", "tf-pose-estimation_b200"))
__path__ = [_real]
__file__ = _os.path.join(_real, "__init__.py")
with open(__file__, "r") as _f:
    exec(compile(_f.read(), __file__, "exec"))
del _f
