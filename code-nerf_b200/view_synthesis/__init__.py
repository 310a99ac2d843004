"""Drop-in `view_synthesis` package: the reference's module paths and call signatures for the volumetric-rendering
hot path (`view_synthesis.nerf`, `view_synthesis.models`), backed by libcnrf.so (sm_100a).

Put `code-nerf_b200/` ahead of the reference checkout on PYTHONPATH.  Everything outside the hot path (`cfgnode`,
`datasets`, `utils`: config, file IO, logging, checkpoints) is NOT re-implemented: when a reference checkout is
available (env CODENERF_REFERENCE, default /root/reference) its `view_synthesis` directory is appended to this
package's search path, so `view_synthesis.utils` etc. resolve to the reference's own unmodified files while
`view_synthesis.nerf` / `view_synthesis.models` resolve here.  train.py / eval.py then run unchanged.
"""
import os as _os

_ref = _os.path.join(_os.environ.get("CODENERF_REFERENCE", "/root/reference"), "view_synthesis")
if _os.path.isdir(_ref) and _ref not in __path__:
    __path__.append(_ref)
