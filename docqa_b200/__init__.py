"""Import alias for `docqa-ms-clinical-document-qa-assistant-llm-microservices-_b200/`.

The contract names the package directory after the reference repository, which is not an
importable identifier; this alias package points its search path at that directory and
executes its __init__, so `import docqa_b200` and `docqa_b200.flat` etc. just work.
"""
import os as _os

_REAL = _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))),
                      "docqa-ms-clinical-document-qa-assistant-llm-microservices-_b200")
__path__.insert(0, _REAL)
with open(_os.path.join(_REAL, "__init__.py")) as _f:
    exec(compile(_f.read(), _os.path.join(_REAL, "__init__.py"), "exec"))
del _f
