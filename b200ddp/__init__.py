"""Alias package: `import b200ddp` resolves to the code under
`distributed-machine-learning-experiment-document_b200/` (whose name is not a
Python identifier)."""
import os as _os

__path__.insert(0, _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))),
                                 "distributed-machine-learning-experiment-document_b200"))
_init = _os.path.join(__path__[0], "__init__.py")
with open(_init) as _f:
    exec(compile(_f.read(), _init, "exec"))
del _os, _f, _init
