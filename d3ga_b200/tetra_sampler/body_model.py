"""Import surface of the reference's `tetra_sampler.body_model`: `SMPLlayer` (lib/smplman.py:9, constructed at :68).

The SMPL / SMPL-X body model itself is outside this repository's scope (it needs the licensed model
files and is pure PyTorch off the render path; what this library accelerates is the skinning of the cage vertices by its
outputs, `d3ga_b200.lbsmodel.smpl_deform`).  So that `lib/smplman.py` still imports when this package stands in for
`tetra_sampler`, `SMPLlayer` here resolves to the reference's own class when its package can be found --

    D3GA_REFERENCE_TETRA_SAMPLER=/path/to/D3GA/submodules/tetra-sampler/tetra_sampler

-- and otherwise to a placeholder whose constructor says so.  Nothing is copied from the reference.
"""
from __future__ import annotations

import importlib.util
import os
import sys

_REF_NAME = "_d3ga_reference_tetra_sampler"


def _load_reference_class():
    root = os.environ.get("D3GA_REFERENCE_TETRA_SAMPLER")
    if not root or not os.path.isfile(os.path.join(root, "body_model.py")):
        return None
    if _REF_NAME not in sys.modules:
        # the reference's body_model does `from .lbs import ...`: give it a private parent package to be relative to
        pkg = importlib.util.module_from_spec(importlib.util.spec_from_loader(_REF_NAME, loader=None, is_package=True))
        pkg.__path__ = [root]
        sys.modules[_REF_NAME] = pkg
    spec = importlib.util.spec_from_file_location(_REF_NAME + ".body_model", os.path.join(root, "body_model.py"))
    mod = importlib.util.module_from_spec(spec)
    sys.modules[spec.name] = mod
    spec.loader.exec_module(mod)
    return mod.SMPLlayer


try:
    SMPLlayer = _load_reference_class()
except Exception as e:  # the reference's own imports (chumpy, model files ...) may be missing
    SMPLlayer = None
    _why = repr(e)
else:
    _why = "D3GA_REFERENCE_TETRA_SAMPLER is not set (or holds no body_model.py)"

if SMPLlayer is None:

    class SMPLlayer:  # noqa: F811
        """Placeholder: the SMPL body model is not part of d3ga-b200 (see the module docstring)."""

        def __init__(self, *args, **kwargs):
            raise ImportError(
                "tetra_sampler.body_model.SMPLlayer is the reference's pure-PyTorch SMPL layer and is not provided by "
                "d3ga-b200; point D3GA_REFERENCE_TETRA_SAMPLER at the reference's tetra_sampler package to use it "
                f"({_why}).")
