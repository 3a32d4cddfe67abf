"""Error codes of the LOWESS path (src/curvefit_core.f90:34-40) plus the one new
code the CUDA library adds (include/curvefit_cuda.h)."""

CF_OUT_OF_MEMORY_ERROR = 105      # curvefit_core.f90:36
CF_INVALID_INPUT_ERROR = 201      # curvefit_core.f90:40
CF_ARRAY_SIZE_ERROR = 202         # curvefit_core.f90:34
CF_NO_DATA_DEFINED_ERROR = 300    # curvefit_core.f90:38
CF_DEVICE_ERROR = 303             # new: CUDA failure (300-302 are taken, :38,43,54)


class CurvefitError(RuntimeError):
    """Raised where the reference's default error object would terminate the
    process (FERROR ferror.f90:164,333: exit(flag) unless disabled)."""

    def __init__(self, fcn: str, msg: str, flag: int):
        super().__init__(f"{fcn}: {msg} (code {flag})")
        self.fcn = fcn
        self.msg = msg
        self.flag = flag
