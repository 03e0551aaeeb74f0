"""semantic-slam_b200 — B200-native TSDF volumetric fusion for semantic-slam's integration
hot path.  The product is semantic-slam_b200/libtsdf_b200.so (CUDA, sm_100a) behind the C ABI
in include/tsdf_b200.h; this package is the thin host-side mirror of the reference's
TSDF / TSDFfusion surface over that ABI.  Import with
`importlib.import_module("semantic-slam_b200")` (the directory name has a hyphen).
"""
from . import capi, synth, volume, dist  # noqa: F401
from .capi import (SEM_TSDFCU, SEM_FUSIONPY, PRED_NONZERO, PRED_ABS_LT, TsdfError, load)  # noqa: F401
from .volume import (Volume, Group, PinnedPool, Batch, TSDF, TSDFVolume, make_config, config_from_bounds, default_config,
                     slab_range, device_count, meshwrite)  # noqa: F401
