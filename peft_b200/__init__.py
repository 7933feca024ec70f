"""Import alias: `import peft_b200` loads the package that lives in the (non-identifier) directory
`parameter-efficient-fine-tuning-for-vision-transformers_b200/` at the repo root."""
import os as _os

_here = _os.path.dirname(_os.path.abspath(__file__))
_real = _os.path.join(_os.path.dirname(_here), "parameter-efficient-fine-tuning-for-vision-transformers_b200")
__path__ = [_real]
with open(_os.path.join(_real, "__init__.py")) as _f:
    exec(compile(_f.read(), _os.path.join(_real, "__init__.py"), "exec"))
