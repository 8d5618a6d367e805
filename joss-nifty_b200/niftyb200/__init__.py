"""niftyb200 -- B200-native correlated-field apply and metric-CG hot path of NIFTy.

Host side (Python) mirrors the reference's interface for this path
(``nifty8.ducc_dispatch``, ``HartleyOperator`` / ``HarmonicTransformOperator``,
``CorrelatedFieldMaker`` / ``SimpleCorrelatedField``, ``ConjugateGradient``, re-style
``cg`` / ``static_cg``); all arithmetic runs in hand-written sm_100a CUDA kernels behind
the C ABI of ``include/niftyb200.h``.  There is no CPU fallback.
"""

from . import config  # noqa: F401
from .transforms import Plan, hartley, fftn, ifftn, vdot, hartley_device  # noqa: F401
from .ducc_dispatch import set_nthreads, nthreads, install  # noqa: F401
from .geometry import RGGrid  # noqa: F401
from .correlated_field import CorrelatedFieldMaker, CorrelatedField, SimpleCorrelatedField  # noqa: F401
from .metric import CFMetric, SampledCFMetric  # noqa: F401
from .cg import (ConjugateGradient, QuadraticEnergy, GradientNormController,  # noqa: F401
                 AbsDeltaEnergyController, cg, static_cg, CGResults)
from .latent import LatentLayout, LatentMetric, draw_linear_residual  # noqa: F401
from .optimize import newton_cg, OptimizeResults  # noqa: F401
from .evi import GaussianCFModel, nonlinearly_update_residual  # noqa: F401
from .los import SamplingCartesianGridLOS, LOSMetric  # noqa: F401
from .distributed import SlabCorrelatedField, SlabHartley, SymmetricSplit  # noqa: F401
from .operators import FFTOperator, HartleyOperator, HarmonicTransformOperator, RGSpace  # noqa: F401

__version__ = "0.1.0"
