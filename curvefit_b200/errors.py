"""Mirror of FERROR's `errors` type as far as the LOWESS path uses it
(FERROR/src/ferror.f90:150-224 type, :282-334 report_error).

The reference's routines take an optional `err` object.  report_error prints,
appends to a log file, stores the flag and message, runs an optional callback
and -- unless exit-on-error was switched off -- terminates the process with
exit code = flag.  A Python host must not kill the interpreter, so "terminate"
becomes raising CurvefitError; everything else is kept: with
set_exit_on_error(False) the call returns and the caller can poll
has_error_occurred()/get_error_flag() exactly as with the Fortran object.
"""
from __future__ import annotations

import datetime
from typing import Callable, Optional

from .core import CurvefitError


class errors:
    def __init__(self):
        self._fname = "error_log.txt"          # ferror.f90:155
        self._found_error = False
        self._flag = 0
        self._exit = True                      # ferror.f90:164
        self._suppress = False
        self._log = False                      # python mirror: no log file unless asked
        self._msg = ""
        self._fcn = ""
        self._callback: Optional[Callable] = None

    # -- configuration ------------------------------------------------------
    def get_log_filename(self) -> str:
        return self._fname

    def set_log_filename(self, name: str, enable: bool = True) -> None:
        self._fname = name
        self._log = enable

    def get_exit_on_error(self) -> bool:
        return self._exit

    def set_exit_on_error(self, x: bool) -> None:
        self._exit = bool(x)

    def get_suppress_printing(self) -> bool:
        return self._suppress

    def set_suppress_printing(self, x: bool) -> None:
        self._suppress = bool(x)

    def set_clean_up_routine(self, fn: Optional[Callable]) -> None:
        self._callback = fn

    # -- reporting (ferror.f90:282-334) ----------------------------------------
    def report_error(self, fcn: str, msg: str, flag: int, obj=None) -> None:
        self._found_error = True
        self._flag = int(flag)
        self._fcn = fcn
        self._msg = msg
        text = (f"\n***** ERROR *****\n{datetime.datetime.now():%Y-%m-%d %H:%M:%S}\n"
                f"Function: {fcn}\nError Flag: {flag}\nMessage:\n{msg}\n")
        if not self._suppress:
            print(text)
        if self._log:
            with open(self._fname, "a") as fh:
                fh.write(text)
        if self._callback is not None:
            self._callback(self, obj)
        if self._exit:
            raise CurvefitError(fcn, msg, int(flag))

    # -- queries ------------------------------------------------------------------
    def has_error_occurred(self) -> bool:
        return self._found_error

    def get_error_flag(self) -> int:
        return self._flag

    def get_error_message(self) -> str:
        return self._msg

    def get_error_fcn_name(self) -> str:
        return self._fcn

    def reset_error_status(self) -> None:
        self._found_error = False
        self._flag = 0
        self._msg = ""
        self._fcn = ""
