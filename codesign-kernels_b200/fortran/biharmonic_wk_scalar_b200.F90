!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!
!! B200 KERNEL ROUTINES: thin ISO_C_BINDING shim onto libbwk_b200.so (include/bwk_b200.h)
!!
!! Drop-in for the reference driver atmosphere/biharmonic_wk_kernel.F90: add
!!   use biharmonic_wk_scalar_b200, only: bws_b200 => biharmonic_wk_scalar
!! next to the three `use` lines at F90:547-549 and call
!!   call bws_b200(elem,qtens,grads,deriv,nets,nete)
!! exactly like bws_gpu2 at F90:574.  Array layouts are untouched: elem(:) is a contiguous
!! array of all-real(8) derived types (F90:23-27), so c_loc(elem(1)) is what the C side
!! reads as bwk_element_t[nelemd].
!!
!! Shipped as source: the build image has no Fortran compiler.  tests/test_abi.py lints the
!! public argument lists with numpy.f2py.crackfortran.
!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!!
module biharmonic_wk_scalar_b200
  use iso_c_binding
  use kernel_utils
  implicit none
  private

  public :: biharmonic_wk_scalar
  public :: biharmonic_wk_scalar_cpusig
  public :: b200_init
  public :: b200_finalize

  interface
    function bwk_init(nelemd_c, nlev_c, qsize_c) bind(c, name="bwk_init") result(rc)
      use iso_c_binding
      integer(c_int), value :: nelemd_c, nlev_c, qsize_c
      integer(c_int) :: rc
    end function bwk_init
    function bwk_finalize() bind(c, name="bwk_finalize") result(rc)
      use iso_c_binding
      integer(c_int) :: rc
    end function bwk_finalize
    function bwk_biharmonic_wk_scalar(elem_c, qtens_c, grads_c, deriv_c, nets_c, nete_c) &
        bind(c, name="bwk_biharmonic_wk_scalar") result(rc)
      use iso_c_binding
      type(c_ptr), value :: elem_c, qtens_c, grads_c, deriv_c
      integer(c_int), intent(in) :: nets_c, nete_c
      integer(c_int) :: rc
    end function bwk_biharmonic_wk_scalar
    function bwk_biharmonic_wk_scalar_nets(elem_c, qtens_c, deriv_c, nets_c, nete_c) &
        bind(c, name="bwk_biharmonic_wk_scalar_nets") result(rc)
      use iso_c_binding
      type(c_ptr), value :: elem_c, qtens_c, deriv_c
      integer(c_int), intent(in) :: nets_c, nete_c
      integer(c_int) :: rc
    end function bwk_biharmonic_wk_scalar_nets
  end interface

contains

  !! once, after allocate_data(): the analogue of `!$acc enter data pcreate(...)` (F90:553)
  subroutine b200_init()
    implicit none
    if (bwk_init(int(nelemd,c_int), int(nlev,c_int), int(qsize,c_int)) /= 0) error stop "bwk_init failed"
  end subroutine b200_init

  subroutine b200_finalize()
    implicit none
    if (bwk_finalize() /= 0) error stop "bwk_finalize failed"
  end subroutine b200_finalize

  !! same dummy arguments as the push-loop variant, F90:525-531
  subroutine biharmonic_wk_scalar(elem,qtens,grads,deriv,nets,nete)
    implicit none
    type (element_t)     , intent(inout), target :: elem(:)
    real (kind=real_kind), intent(inout), target :: qtens(np,np,nlev,qsize,nelemd)
    real (kind=real_kind), intent(inout), target :: grads(np,np,2,nlev,qsize,nelemd)
    type (derivative_t)  , intent(in   ), target :: deriv
    integer              , intent(in   ) :: nets,nete
    integer(c_int) :: rc
    rc = bwk_biharmonic_wk_scalar(c_loc(elem(1)), c_loc(qtens), c_loc(grads), c_loc(deriv), &
                                  int(nets,c_int), int(nete,c_int))
    if (rc /= 0) error stop "bwk_biharmonic_wk_scalar failed"
  end subroutine biharmonic_wk_scalar

  !! same dummy arguments as the CPU and compiler-inline variants, F90:186-191 / F90:317-323
  subroutine biharmonic_wk_scalar_cpusig(elem,qtens,deriv,nets,nete)
    implicit none
    type (element_t)     , intent(in   ), target :: elem(:)
    integer              , intent(in   ) :: nets
    integer              , intent(in   ) :: nete
    real (kind=real_kind), intent(inout), target :: qtens(np,np,nlev,qsize,nets:nete)
    type (derivative_t)  , intent(in   ), target :: deriv
    integer(c_int) :: rc
    rc = bwk_biharmonic_wk_scalar_nets(c_loc(elem(1)), c_loc(qtens), c_loc(deriv), &
                                       int(nets,c_int), int(nete,c_int))
    if (rc /= 0) error stop "bwk_biharmonic_wk_scalar_nets failed"
  end subroutine biharmonic_wk_scalar_cpusig

end module biharmonic_wk_scalar_b200
