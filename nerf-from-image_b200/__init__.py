"""nfi_b200 -- B200-native (sm_100a) volume-rendering hot path of nerf-from-image.

Import as `nfi_b200` (see the shim package of that name at the repository root).
"""
__version__ = "0.1.0"
