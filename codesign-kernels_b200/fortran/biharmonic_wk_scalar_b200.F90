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
  public :: biharmonic_wk_scalar_present
  public :: advance_hypervis_scalar
  public :: biharmonic_wk_dp3d
  public :: b200_set_topology_gdof
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
    function bwk_set_topology_gdof(nelem_global_c, gdof_c, elem_first_c) bind(c, name="bwk_set_topology_gdof") result(rc)
      use iso_c_binding
      integer(c_int64_t), value :: nelem_global_c, elem_first_c
      type(c_ptr), value :: gdof_c
      integer(c_int) :: rc
    end function bwk_set_topology_gdof
    function bwk_advance_hypervis_scalar(elem_c, qdp_c, dp_c, dpdiss_c, deriv_c, dt2_c, nu_q_c, subcycle_c, limiter_c) &
        bind(c, name="bwk_advance_hypervis_scalar") result(rc)
      use iso_c_binding
      type(c_ptr), value :: elem_c, qdp_c, dp_c, dpdiss_c, deriv_c
      real(c_double), value :: dt2_c, nu_q_c
      integer(c_int), value :: subcycle_c, limiter_c
      integer(c_int) :: rc
    end function bwk_advance_hypervis_scalar
    function bwk_biharmonic_wk_dp3d(elem_c, v2c_c, dptens_c, ttens_c, vtens_c, deriv_c, var_coef1_c, undamprrcart_c) &
        bind(c, name="bwk_biharmonic_wk_dp3d") result(rc)
      use iso_c_binding
      type(c_ptr), value :: elem_c, v2c_c, dptens_c, ttens_c, vtens_c, deriv_c
      integer(c_int), value :: var_coef1_c, undamprrcart_c
      integer(c_int) :: rc
    end function bwk_biharmonic_wk_dp3d
    function bwk_biharmonic_wk_scalar_noalloc(d_elem_c, d_qtens_c, deriv_c, nelemd_c, nlev_c, qsize_c, stream_c) &
        bind(c, name="bwk_biharmonic_wk_scalar_noalloc") result(rc)
      use iso_c_binding
      type(c_ptr), value :: d_elem_c, d_qtens_c, deriv_c
      integer(c_int), value :: nelemd_c, nlev_c, qsize_c
      integer(c_int64_t), value :: stream_c
      integer(c_int) :: rc
    end function bwk_biharmonic_wk_scalar_noalloc
    function cudaDeviceSynchronize() bind(c, name="cudaDeviceSynchronize") result(rc)
      use iso_c_binding
      integer(c_int) :: rc
    end function cudaDeviceSynchronize
  end interface

contains

  !! once, after allocate_data(): the analogue of `!$acc enter data pcreate(...)` (F90:553).
  !! Without arguments the context takes kernel_utils' compile-time nlev and qsize (F90:11-13), which is what
  !! biharmonic_wk_scalar and advance_hypervis_scalar need.  biharmonic_wk_dp3d works on four fields per level
  !! (T, v(1), v(2), dp3d) and needs its own context: call b200_finalize(), then b200_init(qsize_ctx=4).
  subroutine b200_init(nlev_ctx, qsize_ctx)
    implicit none
    integer, intent(in), optional :: nlev_ctx, qsize_ctx
    integer :: nl, nq
    nl = nlev
    nq = qsize
    if (present(nlev_ctx)) nl = nlev_ctx
    if (present(qsize_ctx)) nq = qsize_ctx
    if (bwk_init(int(nelemd,c_int), int(nl,c_int), int(nq,c_int)) /= 0) error stop "bwk_init failed"
  end subroutine b200_init

  subroutine b200_finalize()
    implicit none
    if (bwk_finalize() /= 0) error stop "bwk_finalize failed"
  end subroutine b200_finalize

  !! HOMME global dof ids per GLL point, gdof(np,np,nelemd) (the information edgeVpack/edgeVunpack's maps carry):
  !! needed once before the DSS-containing calls below
  subroutine b200_set_topology_gdof(gdof, nelem_global, elem_first)
    implicit none
    integer(c_int64_t), intent(in), target :: gdof(np,np,nelemd)
    integer(c_int64_t), intent(in) :: nelem_global, elem_first
    if (bwk_set_topology_gdof(nelem_global, c_loc(gdof), elem_first) /= 0) error stop "bwk_set_topology_gdof failed"
  end subroutine b200_set_topology_gdof

  !! The caller of this path in HOMME (prim_advection: advance_hypervis_scalar), kernel-extraction style: the state the
  !! upstream routine reads through elem(ie)%state / elem(ie)%derived is passed as plain arrays.
  !!   Qdp(np,np,nlev,qsize,nelemd) = elem(:)%state%Qdp(:,:,:,:,nt_qdp)          (updated in place)
  !!   dp (np,np,nlev,nelemd)       = derived%dp - dt2*derived%divdp_proj
  !!   dpdiss(np,np,nlev,nelemd)    = derived%dpdiss_ave, or dp0(k) broadcast
  subroutine advance_hypervis_scalar(elem,Qdp,dp,dpdiss,deriv,dt2,nu_q,hypervis_subcycle_q,limiter)
    implicit none
    type (element_t)     , intent(in   ), target :: elem(:)
    real (kind=real_kind), intent(inout), target :: Qdp(np,np,nlev,qsize,nelemd)
    real (kind=real_kind), intent(in   ), target :: dp(np,np,nlev,nelemd)
    real (kind=real_kind), intent(in   ), target :: dpdiss(np,np,nlev,nelemd)
    type (derivative_t)  , intent(in   ), target :: deriv
    real (kind=real_kind), intent(in   ) :: dt2, nu_q
    integer              , intent(in   ) :: hypervis_subcycle_q
    logical              , intent(in   ) :: limiter
    integer(c_int) :: rc
    rc = bwk_advance_hypervis_scalar(c_loc(elem(1)), c_loc(Qdp), c_loc(dp), c_loc(dpdiss), c_loc(deriv), &
                                     real(dt2,c_double), real(nu_q,c_double), int(hypervis_subcycle_q,c_int), &
                                     merge(1_c_int, 0_c_int, limiter))
    if (rc /= 0) error stop "bwk_advance_hypervis_scalar failed"
  end subroutine advance_hypervis_scalar

  !! Upstream HOMME biharmonic_wk_dp3d (viscosity_mod), kernel-extraction style: the state the upstream routine reads through
  !! elem(ie)%state is passed in the output arrays (ttens = T, dptens = dp3d, vtens = v on entry, as qtens is the input
  !! of biharmonic_wk_scalar), elem(ie)%vec_sphere2cart as a plain array.  Needs a context made with b200_init(qsize_ctx=4)
  !! (the library rejects any other qsize for this call) and a topology.
  !!   dptens, ttens (np,np,nlev,nelemd) ; vtens (np,np,2,nlev,nelemd) ; vec_sphere2cart (np,np,3,2,nelemd)
  !!   var_coef1 = .false.: no tensorVisc in the first application (hypervis_scaling > 0 upstream)
  subroutine biharmonic_wk_dp3d(elem,vec_sphere2cart,dptens,ttens,vtens,deriv,var_coef1,undamprrcart)
    implicit none
    type (element_t)     , intent(in   ), target :: elem(:)
    real (kind=real_kind), intent(in   ), target :: vec_sphere2cart(np,np,3,2,nelemd)
    real (kind=real_kind), intent(inout), target :: dptens(np,np,nlev,nelemd)
    real (kind=real_kind), intent(inout), target :: ttens(np,np,nlev,nelemd)
    real (kind=real_kind), intent(inout), target :: vtens(np,np,2,nlev,nelemd)
    type (derivative_t)  , intent(in   ), target :: deriv
    logical              , intent(in   )         :: var_coef1, undamprrcart
    integer(c_int) :: rc
    rc = bwk_biharmonic_wk_dp3d(c_loc(elem(1)), c_loc(vec_sphere2cart), c_loc(dptens), c_loc(ttens), c_loc(vtens), c_loc(deriv), &
                                merge(1_c_int, 0_c_int, var_coef1), merge(1_c_int, 0_c_int, undamprrcart))
    if (rc /= 0) error stop "bwk_biharmonic_wk_dp3d failed"
  end subroutine biharmonic_wk_dp3d

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

  !! For an OpenACC build of the driver (pgiacc=1): elem and qtens are already present on the device (`enter data pcreate`,
  !! F90:553, and `update device`, F90:573), so the B200 kernel can run on those device copies in place -- no transfer, no
  !! allocation, no bwk_init (bwk_biharmonic_wk_scalar_noalloc, include/bwk_b200.h; nelemd <= 8192).  The two kernels are queued
  !! on the default stream and waited for here, so the `update host(qtens)` that follows (F90:575) sees the result.
  subroutine biharmonic_wk_scalar_present(elem,qtens,deriv)
    implicit none
    type (element_t)     , intent(in   ), target :: elem(:)
    real (kind=real_kind), intent(inout), target :: qtens(np,np,nlev,qsize,nelemd)
    type (derivative_t)  , intent(in   ), target :: deriv
    integer(c_int) :: rc
    integer(c_int64_t) :: default_stream
    default_stream = 0
    !$acc host_data use_device(elem, qtens)
    rc = bwk_biharmonic_wk_scalar_noalloc(c_loc(elem(1)), c_loc(qtens), c_loc(deriv), &
                                          int(nelemd,c_int), int(nlev,c_int), int(qsize,c_int), default_stream)
    !$acc end host_data
    if (rc /= 0) error stop "bwk_biharmonic_wk_scalar_noalloc failed"
    if (cudaDeviceSynchronize() /= 0) error stop "cudaDeviceSynchronize failed"
  end subroutine biharmonic_wk_scalar_present

end module biharmonic_wk_scalar_b200
