! ref_driver.f90 -- runs the UNTOUCHED reference (jchristopherson/curvefit) on a binary input
! file and dumps what lowess_smoothing%smooth returns, so that the C restatement in
! oracle/lowess_oracle.c can be compared with it bit for bit (tests/test_oracle.py,
! test_oracle_matches_reference_dumps).  Built only where a Fortran compiler exists
! (oracle/ref/build_ref.sh); nothing in the product or in bench.py depends on it.
!
!   ref_driver in.bin out.bin
!   in.bin : int32 n, real64 f, int32 srt (0/1), real64 x(n), real64 y(n)      (stream access)
!   out.bin: int32 n, real64 xs(n) (sorted x), real64 ysrt(n) (co-permuted y), real64 ys(n),
!            real64 res(n)
! nsteps = 2 and delta = 0 are fixed by the reference's ls_smooth (curvefit_regression.f90:866).
program ref_driver
    use iso_fortran_env, only : int32, real64
    use curvefit_regression, only : lowess_smoothing
    implicit none
    type(lowess_smoothing) :: fit
    character(len = 1024) :: fin, fout
    integer(int32) :: n, srt, i, u
    real(real64) :: f
    real(real64), allocatable :: x(:), y(:), ys(:), res(:), xs(:), ysrt(:)

    call get_command_argument(1, fin)
    call get_command_argument(2, fout)
    open(newunit = u, file = trim(fin), access = "stream", form = "unformatted", status = "old")
    read(u) n, f, srt
    allocate(x(n), y(n), res(n), xs(n), ysrt(n))
    read(u) x, y
    close(u)

    call fit%initialize(x, y, srt /= 0)
    ys = fit%smooth(f)
    call fit%get_residuals(res)
    do i = 1, n
        xs(i) = fit%get_x(i)
        ysrt(i) = fit%get_y(i)
    end do

    open(newunit = u, file = trim(fout), access = "stream", form = "unformatted", status = "replace")
    write(u) n, xs, ysrt, ys, res
    close(u)
end program ref_driver
