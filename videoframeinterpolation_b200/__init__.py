"""vfi_b200: B200-native flow warping / deformable sampling behind the reference's Python signatures.

Module layout mirrors the reference's hot-path files (modules/warp.py, softsplat.py, dcnv2.py,
deformable_attn.py, madat.py, models/DAT.py); all arithmetic of the path runs in libvfi_b200.so.
"""
from .warp import bwarp, fwarp, fwarp_using_two_frames, fwarp_mframes
from .softsplat import softsplat, softsplat_func
from .dcnv2 import deform_conv2d, DeformableConv2d, DeformableConv2dGivenOffset
from .deformable_attn import Attn, CrossDeformableAttentionBlockwFlow
from .madat import DeformAttnwMotion, DATwithMotionEstimation
from .dat import DATwConstantnCv1, DCNInterFeatBuilderwithT, InputPadder, default_args
from . import dcndat
from .dcndat import DCNDATv1, dcndat_default_args
from .graph import GraphedInference, PipelinedInference
from .metrics import calculate_psnr, calculate_ssim, frame_metrics
from .checkpoint import load_reference_checkpoint, save_reference_checkpoint, strip_ddp_prefix

FunctionSoftsplat = softsplat   # name used by BASELINE.json's north_star for the same four-argument call

__version__ = "0.1.0"
