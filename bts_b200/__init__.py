"""bts_b200 -- B200-native (sm_100a) implementation of the BTS LPG decoder hot path.

Scope: reduction_1x1 + local_planar_guidance of minghanz/bts (pytorch/bts.py:97-173) and the fused
decoder stage built from them. Everything else in BTS stays stock PyTorch. See DESIGN.md.
"""
from .modules import reduction_1x1, local_planar_guidance, silog_loss, upconv, fused_stage, graft, chain_channels  # noqa: F401
from . import functional  # noqa: F401

__all__ = ["reduction_1x1", "local_planar_guidance", "silog_loss", "upconv", "fused_stage", "graft", "chain_channels", "functional"]
