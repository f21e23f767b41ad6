! $Id$
!
!  GPU = gpu_b200 : drop-in replacement for src/nogpu.f90 / src/gpu_astaroth.f90.
!  Same module name and public list (src/gpu.h); every routine forwards to the
!  C ABI of libpcb200.so (include/pcb200.h) through ISO_C_BINDING, so no symbol
!  mangling headers (cparam_c.h, cdata_c.h, ...) and no Astaroth are needed.
!
!  Select with   GPU = gpu_b200   in src/Makefile.local (see INTEGRATION.md).
!  NOTE: this image has no Fortran compiler; the file is written against the
!  interfaces of the reference (cdata.f90, cparam.f90, the module *_run_pars)
!  and is exercised here only through its C side (tests drive the same entry
!  points from Python).
!
!** AUTOMATIC CPARAM.INC GENERATION ****************************
! Declare (for generation of cparam.inc) the number of f array
! variables and auxiliary variables added by this module
!
! CPARAM logical, parameter :: lgpu = .true.
!
!***************************************************************
module GPU
!
  use Cdata
  use General, only: keep_compiler_quiet
  use Messages, only: fatal_error
  use, intrinsic :: iso_c_binding
!
  implicit none
!
  include 'gpu.h'
!
  integer, parameter :: PCB200_MAXVAR=8, PCB200_ABI_VERSION=1
!
!  mirror of struct pcb200_params (include/pcb200.h) -- keep the order identical
!
  type, bind(C) :: pcb200_params
    integer(c_int32_t) :: abi_version
    integer(c_int32_t) :: nx, ny, nz
    integer(c_int32_t) :: nprocx, nprocy, nprocz
    integer(c_int32_t) :: ipx, ipy, ipz
    integer(c_int32_t) :: lperi(3)
    integer(c_int32_t) :: mvar, mfarray
    integer(c_int32_t) :: iux, iuy, iuz, ilnrho, iss, iax, iay, iaz
    integer(c_int32_t) :: lhydro, ldensity, lentropy, lmagnetic, lgravz, lforcing
    integer(c_int32_t) :: lupw_uu, lupw_lnrho, lupw_ss
    integer(c_int32_t) :: lbidiagonal_derij
    integer(c_int32_t) :: iheatcond, lcooling, lviscosity_heat, lohmic_heat
    integer(c_int32_t) :: ldt
    integer(c_int32_t) :: reserved_i(7)
    integer(c_int32_t) :: bcx(PCB200_MAXVAR,2), bcy(PCB200_MAXVAR,2), bcz(PCB200_MAXVAR,2)
    real(c_double) :: dx, dy, dz
    real(c_double) :: dx_1, dy_1, dz_1
    real(c_double) :: dz2_bound(7)
    real(c_double) :: xyz0(3), Lxyz(3)
    real(c_double) :: x0_loc, y0_loc, z0_loc
    real(c_double) :: cdt, cdtv, cdts, dtmax, dt_fixed
    real(c_double) :: alpha_ts(3), beta_ts(3)
    real(c_double) :: nu, eta, mu0
    real(c_double) :: cs20, gamma, cp, lnrho0
    real(c_double) :: Omega, gravz, chi
    real(c_double) :: hcond0, hcond1, hcond2
    real(c_double) :: z1, z2, widthss
    real(c_double) :: cool, wcool, cs2cool
    real(c_double) :: ztop, Fbot, cs2top
    real(c_double) :: reserved_d(8)
  endtype pcb200_params
!
  type, bind(C) :: pcb200_particle_params
    integer(c_int32_t) :: abi_version
    integer(c_int32_t) :: npar_loc, mpar_loc
    integer(c_int32_t) :: ldragforce_dust_par, ldragforce_gas_par, lcoriolis_force_par
    integer(c_int32_t) :: ldt_adv_par, lcheck_exact_frontier
    integer(c_int32_t) :: reserved_i(8)
    real(c_double) :: tausp, rhop_swarm, cdtp, cdtp_drag
    real(c_double) :: reserved_d(8)
  endtype pcb200_particle_params
!
  type, bind(C) :: pcb200_forcing
    real(c_double) :: coef1(3), coef2(3), fda(3), kk(3)
    real(c_double) :: phase, fact
  endtype pcb200_forcing
!
  interface
    integer(c_int) function pcb200_init(local_rank) bind(C,name='pcb200_init')
      import :: c_int
      integer(c_int), value :: local_rank
    endfunction
    integer(c_int) function pcb200_initialize(par) bind(C,name='pcb200_initialize')
      import :: c_int, pcb200_params
      type(pcb200_params), intent(in) :: par
    endfunction
    integer(c_int) function pcb200_register(f) bind(C,name='pcb200_register')
      import :: c_int, c_double
      real(c_double), dimension(*) :: f
    endfunction
    integer(c_int) function pcb200_substep(isubstep,full,early_finalize,ldiagnos,dt) bind(C,name='pcb200_substep')
      import :: c_int, c_double
      integer(c_int), value :: isubstep, full, early_finalize, ldiagnos
      real(c_double), intent(inout) :: dt
    endfunction
    integer(c_int) function pcb200_copy_farray(f) bind(C,name='pcb200_copy_farray')
      import :: c_int, c_double
      real(c_double), dimension(*) :: f
    endfunction
    integer(c_int) function pcb200_set_forcing(frc) bind(C,name='pcb200_set_forcing')
      import :: c_int, pcb200_forcing
      type(pcb200_forcing), intent(in) :: frc
    endfunction
    integer(c_int) function pcb200_set_zprofile(which,prof,nz) bind(C,name='pcb200_set_zprofile')
      import :: c_int, c_double
      integer(c_int), value :: which, nz
      real(c_double), dimension(*), intent(in) :: prof
    endfunction
    integer(c_int) function pcb200_get_diagnostics(out,n) bind(C,name='pcb200_get_diagnostics')
      import :: c_int, c_double
      real(c_double), dimension(*) :: out
      integer(c_int), value :: n
    endfunction
    integer(c_int) function pcb200_nccl_unique_id(id) bind(C,name='pcb200_nccl_unique_id')
      import :: c_int, c_char
      character(kind=c_char), dimension(128) :: id
    endfunction
    integer(c_int) function pcb200_comm_init(id,rank,nranks) bind(C,name='pcb200_comm_init')
      import :: c_int, c_char
      character(kind=c_char), dimension(128), intent(in) :: id
      integer(c_int), value :: rank, nranks
    endfunction
    integer(c_int) function pcb200_particles_initialize(pp) bind(C,name='pcb200_particles_initialize')
      import :: c_int, pcb200_particle_params
      type(pcb200_particle_params), intent(in) :: pp
    endfunction
    integer(c_int) function pcb200_particles_load(fp,ipar) bind(C,name='pcb200_particles_load')
      import :: c_int, c_double, c_int32_t
      real(c_double), dimension(*), intent(in) :: fp
      integer(c_int32_t), dimension(*), intent(in) :: ipar
    endfunction
    integer(c_int) function pcb200_particles_store(fp,dfp,ipar,ineargrid,k1_imn,k2_imn) bind(C,name='pcb200_particles_store')
      import :: c_int, c_double, c_int32_t
      real(c_double), dimension(*) :: fp, dfp
      integer(c_int32_t), dimension(*) :: ipar, ineargrid, k1_imn, k2_imn
    endfunction
    integer(c_int) function pcb200_particles_get_aux(np,rhop) bind(C,name='pcb200_particles_get_aux')
      import :: c_int, c_double
      real(c_double), dimension(*) :: np, rhop
    endfunction
    integer(c_int) function pcb200_finalize() bind(C,name='pcb200_finalize')
      import :: c_int
    endfunction
    function pcb200_last_error() bind(C,name='pcb200_last_error') result(p)
      import :: c_ptr
      type(c_ptr) :: p
    endfunction
  endinterface
!
  logical, save :: lvery_first=.true.
  real(c_double), dimension(16), save :: gpu_diag
  real, dimension(:,:,:,:), pointer, save :: f_registered => null()   ! the array handed over by register_GPU
!
contains
!***********************************************************************
    subroutine check(rc,where)
!
!  Non-zero status of the C ABI -> fatal_error (the reference C glue ignored
!  AcResult, astaroth/gpu_astaroth.cc:239,341).
!
      integer(c_int), intent(in) :: rc
      character(len=*), intent(in) :: where
      character(kind=c_char), dimension(:), pointer :: msg
      character(len=256) :: str
      integer :: i
!
      if (rc==0) return
      str=''
      call c_f_pointer(pcb200_last_error(),msg,(/256/))
      do i=1,256
        if (msg(i)==c_null_char) exit
        str(i:i)=msg(i)
      enddo
      call fatal_error(where,'libpcb200: '//trim(str))
!
    endsubroutine check
!***********************************************************************
    integer function bc_code(bc)
!
!  run.in boundary-condition string -> PCB200_BC_* (include/pcb200.h)
!
      character(len=*), intent(in) :: bc
!
      select case (trim(bc))
      case ('p');  bc_code=0
      case ('s');  bc_code=1
      case ('a');  bc_code=2
      case ('a2'); bc_code=3
      case ('c1'); bc_code=4
      case ('cT'); bc_code=5
      case default
        bc_code=-1
        call fatal_error('gpu_b200','No GPU implementation for boundary condition "'//trim(bc)//'"')
      endselect
!
    endfunction bc_code
!***********************************************************************
    subroutine gpu_init
!
!  run.f90:125, after mpicomm_init: one rank <-> one B200.  The NCCL
!  communicator is bootstrapped over the host MPI (id broadcast from root).
!
      use Mpicomm, only: mpibcast_char, MPI_COMM_WORLD
!
      character(kind=c_char), dimension(128) :: id
      character(len=128) :: idstr
      integer :: i, local_rank
!
      id=c_null_char
      local_rank=mod(iproc,8)        ! ranks are placed node-contiguously by mpirun
      call check(pcb200_init(int(local_rank,c_int)),'gpu_init')
      if (ncpus>1) then
        if (lroot) call check(pcb200_nccl_unique_id(id),'gpu_init')
        do i=1,128; idstr(i:i)=id(i); enddo
        call mpibcast_char(idstr,0,MPI_COMM_WORLD)
        do i=1,128; id(i)=idstr(i:i); enddo
      endif
      call check(pcb200_comm_init(id,int(iproc,c_int),int(ncpus,c_int)),'gpu_init')
!
    endsubroutine gpu_init
!***********************************************************************
    subroutine initialize_GPU
!
!  register.f90:375 (end of initialize_modules; again on RELOAD, run.f90:571).
!  Refuse module combinations the kernels do not cover (same policy as
!  gpu_astaroth.f90:32-71, shorter list: entropy and gravity ARE covered).
!
!  Only names the reference exports are imported (eos.h:5,50,54, hydro.h:30, density.h:24, gravity.h:18-20);
!  Omega, alpha_ts, beta_ts, mu0, cdt ... come from Cdata.  Module-private run parameters arrive through
!  the reference's own route, pushpars2c (viscosity.f90:2951, magnetic.f90:10646, entropy.f90:8235), which
!  hands out addresses; fortran/host.patch adds the Energy and Magnetic entries the device needs and turns
!  the integer placeholders of noentropy.f90/nomagnetic.f90 into empty routines.
!
      use EquationOfState, only: cs20, gamma, lnrho0, get_cp1, cs2top
      use Hydro, only: lupw_uu
      use Density, only: lupw_lnrho
      use Viscosity, only: viscosity_pars => pushpars2c
      use Magnetic, only: magnetic_pars => pushpars2c
      use Energy, only: energy_pars => pushpars2c
      use Gravity, only: gravz, z1, z2, gravz_profile
!
      type(pcb200_params) :: par
      character(LEN=512) :: str
      real :: cp1
      real(c_double), dimension(nz) :: prof
      integer :: j, k
      integer(KIND=ikind8), dimension(3)  :: p_visc      ! nu, zeta, idiag_dtnu
      integer(KIND=ikind8), dimension(6)  :: p_mag       ! eta, lohmic_heat, idiag_jrms, idiag_jmax, idiag_abm, idiag_jbm
      integer(KIND=ikind8), dimension(18) :: p_ene       ! see fortran/host.patch
      real :: nu, eta, chi, hcond0, hcond1, hcond2, widthss, cool, wcool, cs2cool, Fbot
      logical :: lohmic_heat, lupw_ss, lviscosity_heat, lheatc_Kprof, lheatc_chiconst
      real, dimension(:), pointer :: hcond_prof, dlnhcond_prof
!
!  RELOAD (run.f90:571) while the device owns the state: bring f and the particles back before the device
!  buffers are re-created, so that the upload that follows (lvery_first) starts from the current state
!
      if (.not.lvery_first .and. associated(f_registered)) then
        call check(pcb200_copy_farray(f_registered),'initialize_GPU')
        if (lparticles) call copy_particles_from_GPU(f_registered)
      endif
!
      nu=0.; eta=0.; chi=0.; hcond0=0.; hcond1=0.; hcond2=0.; widthss=0.; cool=0.; wcool=1.; cs2cool=0.; Fbot=0.
      lohmic_heat=.false.; lupw_ss=.false.; lviscosity_heat=.false.; lheatc_Kprof=.false.; lheatc_chiconst=.false.
      nullify(hcond_prof); nullify(dlnhcond_prof)
      call viscosity_pars(p_visc); nu=real_at(p_visc(1))
      if (lmagnetic) then
        call magnetic_pars(p_mag); eta=real_at(p_mag(1)); lohmic_heat=logical_at(p_mag(2))
      endif
      if (lentropy) then
        call energy_pars(p_ene)
        chi=real_at(p_ene(1)); hcond0=real_at(p_ene(2)); hcond1=real_at(p_ene(3)); hcond2=real_at(p_ene(4))
        widthss=real_at(p_ene(5)); cool=real_at(p_ene(6)); wcool=real_at(p_ene(7)); cs2cool=real_at(p_ene(8))
        Fbot=real_at(p_ene(9)); lupw_ss=logical_at(p_ene(10)); lviscosity_heat=logical_at(p_ene(11))
        lheatc_Kprof=logical_at(p_ene(12)); lheatc_chiconst=logical_at(p_ene(13))
        if (p_ene(14)/=0) then
          call c_f_pointer(transfer(p_ene(14),c_null_ptr),hcond_prof,(/nz/))
          call c_f_pointer(transfer(p_ene(15),c_null_ptr),dlnhcond_prof,(/nz/))
        endif
      endif
!
      str=''
      if (lanelastic) str=trim(str)//', '//'anelastic'
      if (lboussinesq) str=trim(str)//', '//'boussinesq'
      if (ltemperature) str=trim(str)//', '//'temperature'
      if (lshock) str=trim(str)//', '//'shock'
      if (lheatflux) str=trim(str)//', '//'heatflux'
      if (lhyperresistivity_strict) str=trim(str)//', '//'hyperresi_strict'
      if (lhyperviscosity_strict) str=trim(str)//', '//'hypervisc_strict'
      if (lADI) str=trim(str)//', '//'implicit_physics'
      if (llorenz_gauge) str=trim(str)//', '//'lorenz_gauge'
      if (ldustvelocity) str=trim(str)//', '//'dustvelocity'
      if (ldustdensity) str=trim(str)//', '//'dustdensity'
      if (ltestscalar) str=trim(str)//', '//'testscalar'
      if (ltestfield) str=trim(str)//', '//'testfield'
      if (ltestflow) str=trim(str)//', '//'testflow'
      if (linterstellar) str=trim(str)//', '//'interstellar'
      if (lcosmicray) str=trim(str)//', '//'cosmicray'
      if (lcosmicrayflux) str=trim(str)//', '//'cosmicrayflux'
      if (lshear) str=trim(str)//', '//'shear'
      if (lpscalar) str=trim(str)//', '//'pscalar'
      if (lradiation) str=trim(str)//', '//'radiation'
      if (lchemistry) str=trim(str)//', '//'chemistry'
      if (lchiral) str=trim(str)//', '//'chiral'
      if (lneutralvelocity) str=trim(str)//', '//'neutralvelocity'
      if (lneutraldensity) str=trim(str)//', '//'neutraldensity'
      if (lpolymer) str=trim(str)//', '//'polymer'
      if (lpointmasses) str=trim(str)//', '//'pointmasses'
      if (lpoisson) str=trim(str)//', '//'poisson'
      if (lselfgravity) str=trim(str)//', '//'selfgravity'
      if (lsolid_cells) str=trim(str)//', '//'solid_cells'
      if (lspecial) str=trim(str)//', '//'special'
      if (lparticles_radius.or.lparticles_number.or.lparticles_density.or.lparticles_sink.or. &
          lparticles_selfgravity.or.lparticles_blocks) str=trim(str)//', '//'particle sub-modules'
      if (.not.lcartesian_coords) str=trim(str)//', '//'curvilinear coordinates'
      if (.not.all(lequidist)) str=trim(str)//', '//'non-equidistant grid'
      if (itorder/=3) str=trim(str)//', '//'itorder/=3'
      if (lgravz .and. gravz_profile/='const') str=trim(str)//', '//'gravz_profile/=const'
      if (lgravx .or. lgravy) str=trim(str)//', '//'gravity along x or y'
!  Not visible from here (module-private, not handed out by pushpars2c): ivisc, iresistivity, lweyl_gauge,
!  hydro's theta, forcing's iforce.  The device implements ivisc='nu-const', iresistivity='eta-const', the Weyl
!  gauge, theta=0 and iforce='helical' (the settings of the samples this path covers); INTEGRATION.md lists them.
      if (str/='') call fatal_error('initialize_GPU','No GPU implementation for "'//trim(str(3:))//'"')
!
      par%abi_version=PCB200_ABI_VERSION
      par%nx=nx; par%ny=ny; par%nz=nz
      par%nprocx=nprocx; par%nprocy=nprocy; par%nprocz=nprocz
      par%ipx=ipx; par%ipy=ipy; par%ipz=ipz
      par%lperi=merge(1,0,lperi)
      par%mvar=mvar; par%mfarray=mfarray
      par%iux=iux; par%iuy=iuy; par%iuz=iuz; par%ilnrho=ilnrho; par%iss=iss
      par%iax=iax; par%iay=iay; par%iaz=iaz
      par%lhydro=merge(1,0,lhydro); par%ldensity=merge(1,0,ldensity)
      par%lentropy=merge(1,0,lentropy); par%lmagnetic=merge(1,0,lmagnetic)
      par%lgravz=merge(1,0,lgravz); par%lforcing=merge(1,0,lforcing)
      par%lupw_uu=merge(1,0,lupw_uu); par%lupw_lnrho=merge(1,0,lupw_lnrho); par%lupw_ss=merge(1,0,lupw_ss)
      par%lbidiagonal_derij=merge(1,0,lbidiagonal_derij)
      par%iheatcond=0
      if (lheatc_Kprof) par%iheatcond=1
      if (lheatc_chiconst) par%iheatcond=2
      if (lheatc_Kprof.and.lheatc_chiconst) call fatal_error('initialize_GPU','No GPU implementation for two iheatcond at once')
      par%lcooling=merge(1,0,cool/=0.)
      par%lviscosity_heat=merge(1,0,lviscosity_heat)
      par%lohmic_heat=merge(1,0,lohmic_heat)
      par%ldt=merge(1,0,ldt)
      par%reserved_i=0
      do j=1,PCB200_MAXVAR; do k=1,2
        par%bcx(j,k)=0; par%bcy(j,k)=0; par%bcz(j,k)=0
        if (j<=mvar) then
          par%bcx(j,k)=bc_code(bcx12(j,k)); par%bcy(j,k)=bc_code(bcy12(j,k)); par%bcz(j,k)=bc_code(bcz12(j,k))
        endif
      enddo; enddo
      par%dx=dx; par%dy=dy; par%dz=dz
      par%dx_1=dx_1(l1); par%dy_1=dy_1(m1); par%dz_1=dz_1(n1)
      par%dz2_bound=0.
      par%dz2_bound(1:3)=dz2_bound(-3:-1); par%dz2_bound(5:7)=dz2_bound(1:3)
      par%xyz0=xyz0; par%Lxyz=Lxyz
      par%x0_loc=x(l1); par%y0_loc=y(m1); par%z0_loc=z(n1)
      par%cdt=cdt; par%cdtv=cdtv; par%cdts=cdts; par%dtmax=dtmax; par%dt_fixed=dt
      par%alpha_ts=alpha_ts(1:3); par%beta_ts=beta_ts(1:3)
      par%nu=nu; par%eta=eta; par%mu0=mu0
      call get_cp1(cp1)
      par%cs20=cs20; par%gamma=gamma; par%cp=1./cp1; par%lnrho0=lnrho0
      par%Omega=Omega; par%gravz=gravz; par%chi=chi
      par%hcond0=hcond0; par%hcond1=hcond1; par%hcond2=hcond2
      par%z1=z1; par%z2=z2; par%widthss=widthss
      par%cool=cool; par%wcool=wcool; par%cs2cool=cs2cool
      par%ztop=xyz0(3)+Lxyz(3); par%Fbot=Fbot; par%cs2top=cs2top
      par%reserved_d=0.
!
      call check(pcb200_initialize(par),'initialize_GPU')
!
!  1-D z profiles the entropy module keeps on the host
!  (get_gravz_heatcond entropy.f90:7243, get_heat_cool_gravz entropy.f90:6244)
!
      if (lentropy.and.par%iheatcond==1) then
        call check(pcb200_set_zprofile(0_c_int,real(hcond_prof,c_double),int(nz,c_int)),'initialize_GPU')
        call check(pcb200_set_zprofile(1_c_int,real(dlnhcond_prof,c_double),int(nz,c_int)),'initialize_GPU')
      endif
      if (lentropy.and.cool/=0.) then
        prof=exp(-0.5*((xyz0(3)+Lxyz(3)-z(n1:n2))/wcool)**2)
        call check(pcb200_set_zprofile(2_c_int,prof,int(nz,c_int)),'initialize_GPU')
      endif
!  particles: their parameters are complete only after particles_initialize_modules, which run.f90 calls AFTER
!  initialize_modules (run.f90:408,429; RELOAD :571-572: rhop_swarm, mp_swarm are computed there), so
!  pcb200_particles_initialize waits for the first rhs_GPU (lvery_first)
      lvery_first=.true.      ! force a full upload after RELOAD / rsnap
!
    endsubroutine initialize_GPU
!***********************************************************************
    subroutine initialize_particles_GPU
!
!  particles_dust run parameters the device needs (particles_dust.f90:59-123,
!  particles_cdata.f90): constant friction time, TSC particle-mesh scheme.
!  tausp, cdtp, cdtp_drag and the four switches are private to module Particles (particles.h exports
!  routines only): they arrive through pushpars2c, which fortran/host.patch adds to particles_dust.f90
!  (and as an empty routine to noparticles.f90) the way viscosity.f90:2951 does it.
!
      use Particles_cdata
      use Particles, only: particles_pars => pushpars2c
!
      type(pcb200_particle_params) :: pp
      integer(KIND=ikind8), dimension(7) :: p_par
      real :: tausp, cdtp, cdtp_drag
      logical :: ldragforce_dust_par, ldragforce_gas_par, lcoriolis_force_par, ldt_adv_par
!
      call particles_pars(p_par)
      if (p_par(1)==0) call fatal_error('initialize_GPU','particles on the GPU need PARTICLES = particles_dust')
      tausp=real_at(p_par(1)); cdtp=real_at(p_par(2)); cdtp_drag=real_at(p_par(3))
      ldragforce_dust_par=logical_at(p_par(4)); ldragforce_gas_par=logical_at(p_par(5))
      lcoriolis_force_par=logical_at(p_par(6)); ldt_adv_par=logical_at(p_par(7))
      if (.not.lparticlemesh_tsc .or. .not.linterpolate_spline) &
          call fatal_error('initialize_GPU','particles on the GPU need lparticlemesh_tsc=T, linterpolate_spline=T')
      pp%abi_version=PCB200_ABI_VERSION
      pp%npar_loc=npar_loc; pp%mpar_loc=mpar_loc
      pp%ldragforce_dust_par=merge(1,0,ldragforce_dust_par)
      pp%ldragforce_gas_par=merge(1,0,ldragforce_gas_par)
      pp%lcoriolis_force_par=merge(1,0,lcoriolis_force_par)
      pp%ldt_adv_par=merge(1,0,ldt_adv_par)
      pp%lcheck_exact_frontier=merge(1,0,lcheck_exact_frontier)
      pp%reserved_i=0
      pp%tausp=tausp; pp%rhop_swarm=rhop_swarm; pp%cdtp=cdtp; pp%cdtp_drag=cdtp_drag
      pp%reserved_d=0.
      call check(pcb200_particles_initialize(pp),'initialize_particles_GPU')
!
    endsubroutine initialize_particles_GPU
!***********************************************************************
    subroutine register_GPU(f)
!
!  run.f90:223 -- hand over the address of f once (aliased, never freed).
!
      real, dimension(:,:,:,:), intent(IN), target, contiguous :: f     ! contiguous: the address of f itself, no copy-in
!
      call check(pcb200_register(f),'register_GPU')
!  kept by address, like registerGPU of the reference (astaroth/gpu_astaroth.cc:224-234): run.f90's f lives for the
!  whole run; the pointer is only used for the download before a RELOAD re-creates the device buffers
      f_registered => f
!
    endsubroutine register_GPU
!***********************************************************************
    subroutine finalize_GPU
!
      call check(pcb200_finalize(),'finalize_GPU')
!
    endsubroutine finalize_GPU
!***********************************************************************
    subroutine rhs_GPU(f,isubstep,early_finalize)
!
!  equ.f90:343 -- one RK substep on the device.  The host skips df=alpha*df,
!  set_dt, the f update and addforce when lgpu (timestep.f90:119-163,
!  run.f90:729); the device does them.  dt comes back on substep 1 so that
!  cdata:dt and t (timestep.f90:152-190) stay right.
!
      use Forcing, only: forcing_pars_hel
      use Particles_cdata, only: ipar
      use Particles_main, only: fp          ! public through fortran/host.patch (particles_main.h exports routines only)
!
      real, dimension (mx,my,mz,mfarray), intent(INOUT) :: f
      integer,                            intent(IN)    :: isubstep
      logical,                            intent(IN)    :: early_finalize
!
      type(pcb200_forcing) :: frc
      real, dimension(3) :: coef1, coef2, coef3, fda, kk
      real :: phase, fact
      real(c_double) :: dtc
!
!  particles travel with f: uploaded once, then owned by the device
!  (the host skips particles_timestep_first/second and particles_boundconds when lgpu)
      if (lparticles .and. lvery_first) then
        call initialize_particles_GPU
        call check(pcb200_particles_load(fp,ipar),'rhs_GPU')
      endif
      dtc=dt
      call check(pcb200_substep(int(isubstep,c_int),merge(1_c_int,0_c_int,lvery_first), &
                                merge(1_c_int,0_c_int,early_finalize),merge(1_c_int,0_c_int,ldiagnos),dtc),'rhs_GPU')
      if (isubstep==1 .and. ldt) dt=dtc
      if (lforcing .and. isubstep==1) then
!  Coefficients of this step's delta-correlated kick (the random numbers stay on the host; one draw per
!  step).  Drawn AFTER substep 1 has returned the new dt: fact=force/ffnorm*sqrt(dt) (forcing.f90:1741),
!  and the reference calls addforce after time_step (run.f90:683,729), i.e. with the dt of this step.
!  The device applies the kick after substep 3.
        call forcing_pars_hel(coef1,coef2,coef3,fda,kk,phase,fact)
        frc%coef1=coef1; frc%coef2=coef2; frc%fda=fda; frc%kk=kk; frc%phase=phase; frc%fact=fact
        call check(pcb200_set_forcing(frc),'rhs_GPU')
      endif
      if (isubstep==1 .and. ldiagnos) then
        call check(pcb200_get_diagnostics(gpu_diag,16_c_int),'rhs_GPU')
        call diagnostics_to_fname
      endif
      lvery_first=.false.
      call keep_compiler_quiet(f)
!
    endsubroutine rhs_GPU
!***********************************************************************
    subroutine copy_farray_from_GPU(f)
!
!  equ.f90:345, snapshot.f90:76,286,302,669 -- all mvar variables incl. ghosts.
!
      real, dimension (mx,my,mz,mfarray), intent(OUT) :: f
!
      call check(pcb200_copy_farray(f),'copy_farray_from_GPU')
      if (lparticles) call copy_particles_from_GPU(f)
!
    endsubroutine copy_farray_from_GPU
!***********************************************************************
    subroutine copy_particles_from_GPU(f)
!
!  Particle state for wsnap/pvar.dat and the host diagnostics: fp, dfp, ipar,
!  ineargrid, the pencil ranges k1_imn/k2_imn and the aux fields np, rhop.
!
      use Particles_cdata
      use Particles_main, only: fp, dfp, ineargrid      ! public through fortran/host.patch
!
      real, dimension (mx,my,mz,mfarray), intent(INOUT) :: f
!
      call check(pcb200_particles_store(fp,dfp,ipar,ineargrid,k1_imn,k2_imn),'copy_particles_from_GPU')
      where (k1_imn/=0)
        npar_imn=k2_imn-k1_imn+1
      elsewhere
        npar_imn=0
      endwhere
      if (inp/=0 .and. irhop/=0) call check(pcb200_particles_get_aux(f(:,:,:,inp),f(:,:,:,irhop)),'copy_particles_from_GPU')
!
    endsubroutine copy_particles_from_GPU
!***********************************************************************
    subroutine diagnostics_to_fname
!
!  The 0-D diagnostics computed on the device (16 sums and maxima of substep 1, PCB200_DIAG_* order of
!  pcb200.h) go straight into fname, so that print.in/time_series.dat work without the full D2H copy and
!  the host pass of equ.f90:344-347 (fortran/host.patch takes ldiagnos out of that branch; 1-D/2-D averages
!  still use it).  fname gets the raw sum or maximum and itype_name the finishing rule of diagnostic()
!  (diagnostics.f90:643-716), exactly what sum_mn_name/max_mn_name leave behind.  The idiag_* indices are
!  module-private: Hydro, Density and Magnetic hand out addresses through pushdiags2c (hydro.f90:8110,
!  density.f90:4016, magnetic.f90:10657); host.patch appends idiag_oum, idiag_dtu (Hydro), idiag_jrms,
!  idiag_jmax, idiag_abm, idiag_jbm (Magnetic), idiag_ssm, idiag_dtc, idiag_dtchi (Energy) and idiag_dtnu
!  (Viscosity) to the same tables.
!
      use Hydro, only: hydro_diags => pushdiags2c
      use Density, only: density_diags => pushdiags2c
      use Magnetic, only: magnetic_diags => pushdiags2c, magnetic_pars => pushpars2c
      use Energy, only: energy_pars => pushpars2c
      use Viscosity, only: viscosity_pars => pushpars2c
!
      integer(KIND=ikind8), dimension(14) :: p_hyd
      integer(KIND=ikind8), dimension(5)  :: p_den
      integer(KIND=ikind8), dimension(8)  :: p_mag
      integer(KIND=ikind8), dimension(6)  :: p_magp
      integer(KIND=ikind8), dimension(18) :: p_ene
      integer(KIND=ikind8), dimension(3)  :: p_visc
!
      call hydro_diags(p_hyd)
      call put(p_hyd(1),gpu_diag(1),ilabel_sum_sqrt)        ! urms: sum u^2
      call put(p_hyd(5),gpu_diag(2),ilabel_max_sqrt)        ! umax: max u^2
      call put(p_hyd(13),gpu_diag(16),ilabel_sum)           ! oum
      call put(p_hyd(14),gpu_diag(10),ilabel_max_dt)        ! dtu: max advec_uu/cdt
      call density_diags(p_den)
      call put(p_den(1),gpu_diag(3),ilabel_sum)             ! rhom
      call viscosity_pars(p_visc)
      call put(p_visc(3),gpu_diag(11),ilabel_max_dt)        ! dtnu
      if (lmagnetic) then
        call magnetic_diags(p_mag)
        call put(p_mag(1),gpu_diag(5),ilabel_sum_sqrt)      ! brms: sum b^2
        call put(p_mag(2),gpu_diag(6),ilabel_max_sqrt)      ! bmax: max b^2
        call magnetic_pars(p_magp)
        call put(p_magp(3),gpu_diag(7),ilabel_sum_sqrt)     ! jrms
        call put(p_magp(4),gpu_diag(8),ilabel_max_sqrt)     ! jmax
        call put(p_magp(5),gpu_diag(14),ilabel_sum)         ! abm
        call put(p_magp(6),gpu_diag(15),ilabel_sum)         ! jbm
      endif
      if (lentropy) then
        call energy_pars(p_ene)
        call put(p_ene(16),gpu_diag(4),ilabel_sum)          ! ssm
        call put(p_ene(17),gpu_diag(9),ilabel_max_dt)       ! dtc
        call put(p_ene(18),gpu_diag(12),ilabel_max_dt)      ! dtchi
      endif
!
    contains
      subroutine put(addr,val,itype)
        integer(KIND=ikind8), intent(in) :: addr
        real(c_double), intent(in) :: val
        integer, intent(in) :: itype
        integer, pointer :: idiag
        if (addr==0) return
        call c_f_pointer(transfer(addr,c_null_ptr),idiag)
        if (idiag/=0) then
          fname(idiag)=val
          itype_name(idiag)=itype
        endif
      endsubroutine put
    endsubroutine diagnostics_to_fname
!***********************************************************************
    real function real_at(addr)
!
!  value of the real module variable whose address pushpars2c handed out (copy_addr, syscalls)
!
      integer(KIND=ikind8), intent(in) :: addr
      real, pointer :: p
      call c_f_pointer(transfer(addr,c_null_ptr),p)
      real_at=p
    endfunction real_at
!***********************************************************************
    logical function logical_at(addr)
      integer(KIND=ikind8), intent(in) :: addr
      logical, pointer :: p
      call c_f_pointer(transfer(addr,c_null_ptr),p)
      logical_at=p
    endfunction logical_at
!***********************************************************************
endmodule GPU
