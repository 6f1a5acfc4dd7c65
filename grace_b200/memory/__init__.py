"""Error-feedback memories (reference ``grace_dl/dist/memory/*.py``)."""
from grace_b200.memory.memories import (  # noqa: F401
    NoneMemory, ResidualMemory, DgcMemory, EFSignSGDMemory, PowerSGDMemory)
