! ISO_C_BINDING view of include/mpdata_b200.h (libmpdata_b200.so): what a maintainer adds to
! mmf-mpdata-tracer/advect_scalar2D_pushncols_openacc.F90 to run its MPDATA step on a B200.
! The wrapper keeps the argument list of advect_scalar2D_cpu / _openacc_1 / _openacc_2 (A:477, A:247, A:72);
! `adz` and the sizes, host-associated in the reference program (A:6-8, A:29), are passed through.
! Shipped as source: this image has no Fortran compiler (tests lint the argument lists only).
module mpdata_b200_mod
  use iso_c_binding, only: c_int, c_double, c_char, c_ptr
  implicit none
  private
  public :: advect_scalar2D_b200

  interface
    function mpdata_advect_scalar2D(nslices, nx, nz, f, u, w, rho, rhow, adz, flux) bind(c, name='mpdata_advect_scalar2D') result(rc)
      import :: c_int, c_double
      integer(c_int), value :: nslices, nx, nz
      real(c_double), intent(inout) :: f(*)
      real(c_double), intent(in)    :: u(*), w(*), rho(*), rhow(*), adz(*)
      real(c_double), intent(inout) :: flux(*)
      integer(c_int) :: rc
    end function mpdata_advect_scalar2D
    function mpdata_last_error() bind(c, name='mpdata_last_error') result(msg)
      import :: c_ptr
      type(c_ptr) :: msg
    end function mpdata_last_error
  end interface

contains

  subroutine advect_scalar2D_b200(f, u, w, rho, rhow, flux, adz, nslices, nx, nz)
    integer, intent(in) :: nslices, nx, nz
    real(c_double), intent(inout) :: f    (nslices, -2:nx+3, 1, nz-1)
    real(c_double), intent(in   ) :: u    (nslices, -1:nx+3, 1, nz-1)
    real(c_double), intent(in   ) :: w    (nslices, -1:nx+2, 1, nz  )
    real(c_double), intent(in   ) :: rho  (nslices, nz-1)
    real(c_double), intent(in   ) :: rhow (nslices, nz  )
    real(c_double), intent(inout) :: flux (nslices, nz  )
    real(c_double), intent(in   ) :: adz  (nslices, nz-1)
    integer(c_int) :: rc
    rc = mpdata_advect_scalar2D(int(nslices, c_int), int(nx, c_int), int(nz, c_int), f, u, w, rho, rhow, adz, flux)
    if (rc /= 0) error stop 'mpdata_advect_scalar2D failed (see mpdata_last_error)'
  end subroutine advect_scalar2D_b200

end module mpdata_b200_mod
