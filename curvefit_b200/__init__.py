"""curvefit_b200 -- B200-native LOWESS behind curvefit's lowess_smoothing API.

Host-side mirror (Python over ctypes) of the reference's Fortran type
`lowess_smoothing` (src/curvefit_regression.f90:218-247); all numerical work
happens in libcurvefit_cuda.so (hand-written sm_100a CUDA, C ABI declared in
include/curvefit_cuda.h).  There is no CPU fallback: without the built
library or without a CUDA device every operation raises.
"""
from .core import (CF_ARRAY_SIZE_ERROR, CF_DEVICE_ERROR, CF_INVALID_INPUT_ERROR,
                   CF_NO_DATA_DEFINED_ERROR, CF_OUT_OF_MEMORY_ERROR, CurvefitError)
from .errors import errors
from .regression import lowess_smoothing, moving_average
from .statistics import median

__all__ = [
    "lowess_smoothing", "moving_average", "median", "errors", "CurvefitError",
    "CF_ARRAY_SIZE_ERROR", "CF_OUT_OF_MEMORY_ERROR", "CF_NO_DATA_DEFINED_ERROR",
    "CF_INVALID_INPUT_ERROR", "CF_DEVICE_ERROR",
]
