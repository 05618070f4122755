"""Drop-in for the reference's top-level `lbsmodel` package on the cage-vertex skinning path.
`from lbsmodel.body_model import LBSModule` (lib/blueman.py:9, lib/cage_blueman.py) resolves here after
`d3ga_b200.install_dropin()`."""
from . import body_model  # noqa: F401
from .body_model import (LBSModule, LinearBlendSkinning, ParameterTransform, compute_root_transform_cuda,  # noqa: F401
                         load_momentum_cfg, smpl_deform, solve_skeleton_state, states_to_matrix)
