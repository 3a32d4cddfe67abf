! curvefit_regression_cuda.f90
!
! Drop-in replacement for the LOWESS part of curvefit's curvefit_regression
! module (reference src/curvefit_regression.f90:218-247 type lowess_smoothing,
! :740-811 ls_init, :832-868 ls_smooth, :876-939 accessors).  Same type name,
! same components, same type-bound procedure names and argument lists, same
! error reporting through FERROR -- but initialize/smooth call the C ABI of
! libcurvefit_cuda (include/curvefit_cuda.h) instead of the Fortran lowess().
!
! NOT COMPILED IN THE DEVELOPMENT IMAGE (no Fortran compiler there); built by
! CMake only when check_language(Fortran) succeeds.  It needs the reference's
! curvefit_core (error codes) and ferror modules on the include path.
!
! The type keeps only allocatable components, exactly like the reference, so
! intrinsic assignment still deep-copies and there is no finaliser: the C
! entry points used here are the STATELESS ones (host buffers in, host buffers
! out), nothing device-side outlives a call.
module curvefit_regression_cuda
    use, intrinsic :: iso_c_binding
    use, intrinsic :: iso_fortran_env, only : int32, real64
    use curvefit_core, only : CF_ARRAY_SIZE_ERROR, CF_OUT_OF_MEMORY_ERROR, &
        CF_NO_DATA_DEFINED_ERROR, CF_INVALID_INPUT_ERROR
    use ferror, only : errors
    implicit none
    private
    public :: lowess_smoothing
    public :: moving_average
    ! additions that have no counterpart in the reference (nothing above changes meaning):
    public :: lowess_set_evaluator
    public :: lowess_use_all_gpus
    public :: lowess_release_gpus
    public :: CF_METHOD_DIRECT, CF_METHOD_TREE, CF_METHOD_AUTO

    !> Same generic name as the reference (src/curvefit_regression.f90:82-84).
    interface moving_average
        module procedure :: moving_average_1
    end interface

    !> Code libcurvefit_cuda returns for a CUDA failure (no reference equivalent).
    integer(int32), parameter :: CF_DEVICE_ERROR = 303
    !> How a lowest() pass is evaluated (include/curvefit_cuda.h): the O(n ns) sweep in the
    !! reference's summation order (default), the O(n log ns) moment tree (equal to rounding,
    !! <= 1e-12 relative), or the tree whenever ns >= 1024.
    integer(int32), parameter :: CF_METHOD_DIRECT = 0
    integer(int32), parameter :: CF_METHOD_TREE = 1
    integer(int32), parameter :: CF_METHOD_AUTO = 2
    !> Once lowess_use_all_gpus(.true.) has been called, series at least this long are
    !! smoothed on all GPUs of the box (cf_lowess_smooth_multi); shorter ones, and every
    !! series by default, on the current device.
    integer(int32), parameter :: CF_MULTI_GPU_MIN_POINTS = 4000000
    !> A direct lowest() pass costs n * ns (point, element) pairs, about 1.5e12 of them per second
    !! on one B200.  From this many pairs per pass on, smooth() asks for the moment tree for that
    !! call (cf_lowess_smooth_ex with CF_METHOD_TREE: equal to <= 1e-12 relative, contract 1e-10),
    !! so that a series of 1e8 points is 0.1 s and not 22 minutes; below it the process default
    !! applies (the direct sweep unless lowess_set_evaluator changed it).
    real(real64), parameter :: CF_TREE_MIN_PAIRS = 1.0d12
    !> Opt-in switch of the multi-GPU route (process-wide, default off): smooth() behaves
    !! exactly as on a one-GPU box unless the program asks for more.
    logical, save :: m_all_gpus = .false.

    type lowess_smoothing
        private
        real(real64), allocatable, dimension(:) :: m_x
        real(real64), allocatable, dimension(:) :: m_y
        real(real64), allocatable, dimension(:) :: m_weights
        real(real64), allocatable, dimension(:) :: m_residuals
        real(real64) :: m_delta
        logical :: m_init = .false.
    contains
        procedure, public :: initialize => ls_init
        procedure, public :: smooth => ls_smooth
        procedure, public :: get_count => ls_get_num_pts
        procedure, public :: get_x => ls_get_x
        procedure, public :: get_y => ls_get_y
        procedure, public :: get_residuals => ls_get_residual
    end type

    interface
        function cf_sort_xy(x, y, n, xs, ys, perm) bind(C, name = "cf_sort_xy") &
                result(rc)
            import :: c_int32_t, c_double, c_ptr
            real(c_double), intent(in) :: x(*), y(*)
            integer(c_int32_t), value :: n
            real(c_double), intent(out) :: xs(*), ys(*)
            type(c_ptr), value :: perm
            integer(c_int32_t) :: rc
        end function

        function cf_lowess_smooth(x, y, n, f, nsteps, ys, rw, res) &
                bind(C, name = "cf_lowess_smooth") result(rc)
            import :: c_int32_t, c_double
            real(c_double), intent(in) :: x(*), y(*)
            integer(c_int32_t), value :: n
            real(c_double), value :: f
            integer(c_int32_t), value :: nsteps
            real(c_double), intent(out) :: ys(*), rw(*), res(*)
            integer(c_int32_t) :: rc
        end function

        ! cf_lowess_smooth with the evaluator given for this call only (thread-local in the
        ! library; the process default is neither read nor changed)
        function cf_lowess_smooth_ex(x, y, n, f, nsteps, method, ys, rw, res) &
                bind(C, name = "cf_lowess_smooth_ex") result(rc)
            import :: c_int32_t, c_double
            real(c_double), intent(in) :: x(*), y(*)
            integer(c_int32_t), value :: n
            real(c_double), value :: f
            integer(c_int32_t), value :: nsteps
            integer(c_int32_t), value :: method
            real(c_double), intent(out) :: ys(*), rw(*), res(*)
            integer(c_int32_t) :: rc
        end function

        ! ns = max(min(int(f * n), n), 2), src/curvefit_regression.f90:647
        function cf_lowess_span(n, f) bind(C, name = "cf_lowess_span") result(ns)
            import :: c_int32_t, c_double
            integer(c_int32_t), value :: n
            real(c_double), value :: f
            integer(c_int32_t) :: ns
        end function

        ! one series over all (or the listed) GPUs of the box from this one process;
        ! devices = c_null_ptr, ndev = 0: every visible device.  Falls back to one GPU
        ! for short series or a single device (include/curvefit_cuda.h).
        function cf_lowess_smooth_multi(x, y, n, f, nsteps, devices, ndev, ys, rw, res) &
                bind(C, name = "cf_lowess_smooth_multi") result(rc)
            import :: c_int32_t, c_double, c_ptr
            real(c_double), intent(in) :: x(*), y(*)
            integer(c_int32_t), value :: n
            real(c_double), value :: f
            integer(c_int32_t), value :: nsteps
            type(c_ptr), value :: devices
            integer(c_int32_t), value :: ndev
            real(c_double), intent(out) :: ys(*), rw(*), res(*)
            integer(c_int32_t) :: rc
        end function

        ! frees the per-device buffers and the peer mapping cf_lowess_smooth_multi keeps
        ! between calls (about 60 bytes per point on every GPU)
        function cf_lowess_multi_release() bind(C, name = "cf_lowess_multi_release") &
                result(rc)
            import :: c_int32_t
            integer(c_int32_t) :: rc
        end function

        function cf_set_default_method(method) bind(C, name = "cf_set_default_method") &
                result(rc)
            import :: c_int32_t
            integer(c_int32_t), value :: method
            integer(c_int32_t) :: rc
        end function

        function cf_moving_average(x, n, npts) bind(C, name = "cf_moving_average") &
                result(rc)
            import :: c_int32_t, c_double
            real(c_double), intent(inout) :: x(*)
            integer(c_int32_t), value :: n, npts
            integer(c_int32_t) :: rc
        end function

        function cf_last_error(buf, len) bind(C, name = "cf_last_error") result(rc)
            import :: c_int32_t, c_char
            character(kind = c_char), intent(out) :: buf(*)
            integer(c_int32_t), value :: len
            integer(c_int32_t) :: rc
        end function
    end interface

contains
    !> Process-wide choice of the evaluator of every later smooth() (CF_METHOD_*).  A series of
    !! n = 1e8 points with f = 0.05 takes 97 ms per smooth on one B200 with the tree and about
    !! 22 minutes with the direct sweep; the default stays the sweep because it sums in the
    !! reference's order.
    subroutine lowess_set_evaluator(method, err)
        integer(int32), intent(in) :: method
        class(errors), intent(inout), optional, target :: err
        class(errors), pointer :: errmgr
        type(errors), target :: deferr
        integer(int32) :: rc
        if (present(err)) then
            errmgr => err
        else
            errmgr => deferr
        end if
        rc = cf_set_default_method(method)
        if (rc /= 0) call report(errmgr, "lowess_set_evaluator", rc)
    end subroutine

    !> Lets smooth() spread long series (CF_MULTI_GPU_MIN_POINTS or more) over every GPU of
    !! the box.  Off by default: the sharded run keeps buffers for the whole series on every
    !! device between calls, which a program sharing the box with other jobs may not want.
    !! Where the GPUs cannot reach each other's memory, or one of them is full, the library
    !! runs the call on one GPU instead of failing (include/curvefit_cuda.h).
    subroutine lowess_use_all_gpus(flag)
        logical, intent(in) :: flag
        integer(int32) :: rc
        m_all_gpus = flag
        if (.not.flag) rc = cf_lowess_multi_release()
    end subroutine

    !> Frees what the multi-GPU route keeps on the devices between calls.  The type has no
    !! finaliser (allocatable components only, like the reference, so that assignment still
    !! deep-copies): a program that used lowess_use_all_gpus(.true.) calls this when it is
    !! done smoothing, or lowess_use_all_gpus(.false.), which does the same.
    subroutine lowess_release_gpus()
        integer(int32) :: rc
        rc = cf_lowess_multi_release()
    end subroutine

    !> Reports a non-zero libcurvefit_cuda status through the reference's error
    !! object with the reference's function names and messages.
    subroutine report(errmgr, fcn, rc)
        class(errors), intent(inout) :: errmgr
        character(len = *), intent(in) :: fcn
        integer(int32), intent(in) :: rc
        character(kind = c_char) :: cbuf(512)
        character(len = 512) :: msg
        integer(int32) :: i, k
        select case (rc)
        case (CF_ARRAY_SIZE_ERROR)
            call errmgr%report_error(fcn, "Input array sizes must match.", rc)
        case (CF_OUT_OF_MEMORY_ERROR)
            call errmgr%report_error(fcn, "Insufficient memory available.", rc)
        case (CF_NO_DATA_DEFINED_ERROR)
            call errmgr%report_error(fcn, "No data has been defined.", rc)
        case default
            k = cf_last_error(cbuf, 512_int32)
            msg = " "
            do i = 1, min(k, 511)
                msg(i:i) = cbuf(i)
            end do
            call errmgr%report_error(fcn, trim(msg), rc)
        end select
    end subroutine

    ! reference moving_average_1, src/curvefit_regression.f90:427-478 (bit-identical result)
    subroutine moving_average_1(x, npts, err)
        real(real64), intent(inout), dimension(:) :: x
        integer(int32), intent(in) :: npts
        class(errors), intent(inout), optional, target :: err

        integer(int32) :: i, n, flag, rc
        real(real64), allocatable, dimension(:) :: xc
        class(errors), pointer :: errmgr
        type(errors), target :: deferr

        n = size(x)
        if (present(err)) then
            errmgr => err
        else
            errmgr => deferr
        end if
        if (npts < 2 .or. npts > n) then
            call errmgr%report_error("moving_average_1", &
                "The averaging size window must be at least 2, and less " // &
                "than the size of the data set.", CF_INVALID_INPUT_ERROR)
            return
        end if
        ! x is assumed-shape and may be strided: the C entry point takes a contiguous copy
        allocate(xc(n), stat = flag)
        if (flag /= 0) then
            call errmgr%report_error("moving_average_1", &
                "Insufficient memory available.", CF_OUT_OF_MEMORY_ERROR)
            return
        end if
        do concurrent (i = 1:n)
            xc(i) = x(i)
        end do
        rc = cf_moving_average(xc, n, npts)
        if (rc /= 0) then
            call report(errmgr, "moving_average_1", rc)
            return
        end if
        do concurrent (i = 1:n)
            x(i) = xc(i)
        end do
    end subroutine

    ! reference ls_init, src/curvefit_regression.f90:740-811
    subroutine ls_init(this, x, y, srt, err)
        class(lowess_smoothing), intent(inout) :: this
        real(real64), intent(in), dimension(:) :: x, y
        logical, intent(in), optional :: srt
        class(errors), intent(inout), optional, target :: err

        integer(int32) :: i, n, flag, rc
        real(real64), allocatable, dimension(:) :: xc, yc
        class(errors), pointer :: errmgr
        type(errors), target :: deferr
        logical :: sortData

        this%m_init = .false.
        n = size(x)
        sortData = .true.
        if (present(srt)) sortData = srt
        if (present(err)) then
            errmgr => err
        else
            errmgr => deferr
        end if

        if (size(y) /= n) then
            call errmgr%report_error("ls_init", &
                "Input array sizes must match.", CF_ARRAY_SIZE_ERROR)
            return
        end if

        if (allocated(this%m_x)) deallocate(this%m_x)
        if (allocated(this%m_y)) deallocate(this%m_y)
        if (allocated(this%m_weights)) deallocate(this%m_weights)
        if (allocated(this%m_residuals)) deallocate(this%m_residuals)
        allocate(this%m_x(n), stat = flag)
        if (flag == 0) allocate(this%m_y(n), stat = flag)
        if (flag == 0) allocate(this%m_weights(n), stat = flag)
        if (flag == 0) allocate(this%m_residuals(n), stat = flag)
        if (flag == 0 .and. sortData) allocate(xc(n), stat = flag)
        if (flag == 0 .and. sortData) allocate(yc(n), stat = flag)
        if (flag /= 0) then
            call errmgr%report_error("ls_init", &
                "Insufficient memory available.", CF_OUT_OF_MEMORY_ERROR)
            return
        end if

        if (sortData) then
            ! contiguous copies (x, y are assumed-shape and may be strided),
            ! then the device radix sort replaces sort(m_x, indices, .true.)
            do concurrent (i = 1:n)
                xc(i) = x(i)
                yc(i) = y(i)
            end do
            rc = cf_sort_xy(xc, yc, n, this%m_x, this%m_y, c_null_ptr)
            if (rc /= 0) then
                call report(errmgr, "ls_init", rc)
                return
            end if
        else
            do concurrent (i = 1:n)
                this%m_x(i) = x(i)
                this%m_y(i) = y(i)
            end do
        end if

        this%m_delta = 0.0d0
        this%m_init = .true.
    end subroutine

    ! reference ls_smooth, src/curvefit_regression.f90:832-868
    function ls_smooth(this, f, err) result(ys)
        class(lowess_smoothing), intent(inout) :: this
        real(real64), intent(in) :: f
        class(errors), intent(inout), optional, target :: err
        real(real64), allocatable, dimension(:) :: ys

        integer(int32) :: n, flag, rc, ns
        class(errors), pointer :: errmgr
        type(errors), target :: deferr

        if (present(err)) then
            errmgr => err
        else
            errmgr => deferr
        end if
        if (.not.this%m_init) then
            call errmgr%report_error("ls_smooth", &
                "No data has been defined.", CF_NO_DATA_DEFINED_ERROR)
            return
        end if
        n = size(this%m_x)

        allocate(ys(n), stat = flag)
        if (flag /= 0) then
            call errmgr%report_error("ls_smooth", &
                "Insufficient memory available.", CF_OUT_OF_MEMORY_ERROR)
            return
        end if
        ! lowess(m_x, m_y, f, 2, m_delta = 0, ys, m_weights, m_residuals), :866 --
        ! range-sharded over every GPU of the box when the program opted in
        ! (lowess_use_all_gpus) and the series is long enough to be worth it (same
        ! results: per-point sums never cross devices), else on the current device
        if (m_all_gpus .and. n >= CF_MULTI_GPU_MIN_POINTS) then
            rc = cf_lowess_smooth_multi(this%m_x, this%m_y, n, f, 2_int32, &
                c_null_ptr, 0_int32, ys, this%m_weights, this%m_residuals)
        else
            ns = cf_lowess_span(n, f)
            if (real(n, real64) * real(ns, real64) >= CF_TREE_MIN_PAIRS) then
                rc = cf_lowess_smooth_ex(this%m_x, this%m_y, n, f, 2_int32, &
                    CF_METHOD_TREE, ys, this%m_weights, this%m_residuals)
            else
                rc = cf_lowess_smooth(this%m_x, this%m_y, n, f, 2_int32, ys, &
                    this%m_weights, this%m_residuals)
            end if
        end if
        if (rc /= 0) call report(errmgr, "ls_smooth", rc)
    end function

    pure function ls_get_num_pts(this) result(n)
        class(lowess_smoothing), intent(in) :: this
        integer(int32) :: n
        if (this%m_init) then
            n = size(this%m_x)
        else
            n = 0
        end if
    end function

    pure function ls_get_x(this, ind) result(x)
        class(lowess_smoothing), intent(in) :: this
        integer(int32), intent(in) :: ind
        real(real64) :: x
        if (this%m_init) then
            x = this%m_x(ind)
        else
            x = 0.0d0
        end if
    end function

    pure function ls_get_y(this, ind) result(y)
        class(lowess_smoothing), intent(in) :: this
        integer(int32), intent(in) :: ind
        real(real64) :: y
        if (this%m_init) then
            y = this%m_y(ind)
        else
            y = 0.0d0
        end if
    end function

    subroutine ls_get_residual(this, x)
        class(lowess_smoothing), intent(in) :: this
        real(real64), intent(out), dimension(:) :: x
        integer(int32) :: n
        n = min(size(x), this%get_count())
        x(1:n) = this%m_residuals(1:n)
    end subroutine
end module
