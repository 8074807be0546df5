"""Import alias: the product package lives in ``chatterbox-vllm2_b200/`` (a name Python cannot
import directly because of the hyphen); this shim makes it importable as ``chatterbox_vllm2_b200``."""
from pathlib import Path as _Path

_real = _Path(__file__).resolve().parent.parent / "chatterbox-vllm2_b200"
__path__ = [str(_real)]
__file__ = str(_real / "__init__.py")
exec(compile((_real / "__init__.py").read_text(), __file__, "exec"))
