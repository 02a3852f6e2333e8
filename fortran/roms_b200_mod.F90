! roms_b200_mod.F90 -- ISO_C_BINDING interface to libromsb200.so (include/roms_b200.h).
!
! Drop this file and roms_b200_shims.F90 into the reference's src/, remove the bodies of the eleven tile
! subroutines they replace (step2d_FB.F:3-22, pre_step3d4S.F:4-21, prsgrd.F:4-23, rho_eos.F:6-24,
! step3d_uv1.F:5-21, step3d_uv2.F:6-16, step3d_t_ISO.F:20-43, omega.F:4-15, set_depth.F:4-14,190-199,239-250)
! and link with -lromsb200.  main.F is unchanged: it keeps calling step2d, pre_step3d(tile), ... (main.F:397-479).
! No Fortran compiler exists in the image this was written in: the file is checked by tests/test_fortran_shims.py
! (every bind(C) name is declared by the header and exported by the library, argument counts agree).
module roms_b200_mod
  use, intrinsic :: iso_c_binding
  implicit none
  public

  type(c_ptr), save :: roms_h = c_null_ptr               ! opaque roms_ctx* of this MPI rank / GPU

  integer(c_int), parameter :: ROMS_MAXNT = 40

  type, bind(C) :: roms_cfg                              ! mirrors struct roms_cfg, field for field
    integer(c_int) :: Lm, Mm, N, NT
    integer(c_int) :: ew_periodic, ns_periodic
    integer(c_int) :: nonlin_eos
    integer(c_int) :: salinity
    integer(c_int) :: ndtfast
    integer(c_int) :: np_xi, np_eta
    integer(c_int) :: inode, jnode
    real(c_double) :: dt
    real(c_double) :: rho0, g
    real(c_double) :: Tcoef, T0, Scoef, S0
    real(c_double) :: rdrg, rdrg2, Zob
    real(c_double) :: gamma2
    real(c_double) :: theta_s, theta_b, hc
    integer(c_int) :: uv_vis2
    integer(c_int) :: ts_dif2
    real(c_double) :: visc2
    real(c_double) :: tnu2(ROMS_MAXNT)
    integer(c_int) :: obc_west, obc_east, obc_south, obc_north
    real(c_double) :: ubind
    integer(c_int) :: obc_m2, obc_m3, obc_t
  end type roms_cfg

  ! enum roms_field (0-based, order of include/roms_b200.h)
  enum, bind(C)
    enumerator :: ROMS_H = 0, ROMS_PM, ROMS_PN, ROMS_F, ROMS_RMASK
    enumerator :: ROMS_UMASK, ROMS_VMASK, ROMS_DM_U, ROMS_DN_U, ROMS_DM_V, ROMS_DN_V, ROMS_DM_R, ROMS_DN_R
    enumerator :: ROMS_FOMN, ROMS_HINV
    enumerator :: ROMS_ZETA, ROMS_UBAR, ROMS_VBAR
    enumerator :: ROMS_RHOA, ROMS_RHOS, ROMS_RUFRC, ROMS_RVFRC
    enumerator :: ROMS_ZT_AVG1, ROMS_DU_AVG1, ROMS_DV_AVG1, ROMS_DU_AVG2, ROMS_DV_AVG2
    enumerator :: ROMS_DU_AVG_BAK, ROMS_DV_AVG_BAK, ROMS_R_D
    enumerator :: ROMS_SUSTR, ROMS_SVSTR, ROMS_SWFLX
    enumerator :: ROMS_STFLX
    enumerator :: ROMS_U, ROMS_V
    enumerator :: ROMS_T
    enumerator :: ROMS_HZ, ROMS_Z_R
    enumerator :: ROMS_Z_W
    enumerator :: ROMS_FLXU, ROMS_FLXV
    enumerator :: ROMS_WE, ROMS_WI
    enumerator :: ROMS_RHO1, ROMS_QP1
    enumerator :: ROMS_AKV
    enumerator :: ROMS_AKT
    enumerator :: ROMS_RU, ROMS_RV
    enumerator :: ROMS_PMASK, ROMS_DM_P, ROMS_DN_P, ROMS_PMON_U, ROMS_PNOM_V
    enumerator :: ROMS_VISC2_R, ROMS_VISC2_P
    enumerator :: ROMS_DIFF2
    enumerator :: ROMS_RIV_UFLX, ROMS_RIV_VFLX
    enumerator :: ROMS_PTIDE
    enumerator :: ROMS_PIPE_IDX, ROMS_PIPE_FLX
    enumerator :: ROMS_SRFLX
    enumerator :: ROMS_ZETA_WEST, ROMS_ZETA_EAST, ROMS_ZETA_SOUTH, ROMS_ZETA_NORTH
    enumerator :: ROMS_UBAR_WEST, ROMS_UBAR_EAST, ROMS_UBAR_SOUTH, ROMS_UBAR_NORTH
    enumerator :: ROMS_VBAR_WEST, ROMS_VBAR_EAST, ROMS_VBAR_SOUTH, ROMS_VBAR_NORTH
    enumerator :: ROMS_U_WEST, ROMS_U_EAST, ROMS_U_SOUTH, ROMS_U_NORTH
    enumerator :: ROMS_V_WEST, ROMS_V_EAST, ROMS_V_SOUTH, ROMS_V_NORTH
    enumerator :: ROMS_T_WEST, ROMS_T_EAST, ROMS_T_SOUTH, ROMS_T_NORTH
    enumerator :: ROMS_NFIELDS
  end enum

  interface
    ! ---- lifetime, data movement
    integer(c_int) function roms_create(cfg, device, out) bind(C, name='roms_create')
      import; type(roms_cfg), intent(in) :: cfg; integer(c_int), value :: device; type(c_ptr), intent(out) :: out
    end function
    integer(c_int) function roms_destroy(h) bind(C, name='roms_destroy')
      import; type(c_ptr), value :: h
    end function
    integer(c_int) function roms_push(h, field, lvl, itrc, host) bind(C, name='roms_push')
      import; type(c_ptr), value :: h, host; integer(c_int), value :: field, lvl, itrc
    end function
    integer(c_int) function roms_pull(h, field, lvl, itrc, host) bind(C, name='roms_pull')
      import; type(c_ptr), value :: h, host; integer(c_int), value :: field, lvl, itrc
    end function
    integer(c_int) function roms_push_async(h, field, lvl, itrc, host) bind(C, name='roms_push_async')
      import; type(c_ptr), value :: h, host; integer(c_int), value :: field, lvl, itrc
    end function
    integer(c_int) function roms_pull_async(h, field, lvl, itrc, host) bind(C, name='roms_pull_async')
      import; type(c_ptr), value :: h, host; integer(c_int), value :: field, lvl, itrc
    end function
    integer(c_int) function roms_sync(h) bind(C, name='roms_sync')
      import; type(c_ptr), value :: h
    end function
    integer(c_int) function roms_setup_grid(h) bind(C, name='roms_setup_grid')
      import; type(c_ptr), value :: h
    end function
    ! ---- loop-control scalars (scalars.F:28-32): the Fortran driver keeps owning the rotation
    integer(c_int) function roms_set_time_indices(h, iic, iif, nstp, nrhs, nnew, kstp, knew) bind(C, name='roms_set_time_indices')
      import; type(c_ptr), value :: h; integer(c_int), value :: iic, iif, nstp, nrhs, nnew, kstp, knew
    end function
    integer(c_int) function roms_compute_tile_bounds(h, tile, istr, iend, jstr, jend) bind(C, name='roms_compute_tile_bounds')
      import; type(c_ptr), value :: h; integer(c_int), value :: tile; integer(c_int), intent(out) :: istr, iend, jstr, jend
    end function
    ! ---- the tile-subroutine surface, one per reference subroutine (main.F:397-479)
    integer(c_int) function roms_set_depth_c(h, tile) bind(C, name='roms_set_depth')
      import; type(c_ptr), value :: h; integer(c_int), value :: tile
    end function
    integer(c_int) function roms_set_huv_c(h) bind(C, name='roms_set_huv')
      import; type(c_ptr), value :: h
    end function
    integer(c_int) function roms_set_huv1_c(h, tile) bind(C, name='roms_set_huv1')
      import; type(c_ptr), value :: h; integer(c_int), value :: tile
    end function
    integer(c_int) function roms_omega_c(h) bind(C, name='roms_omega')
      import; type(c_ptr), value :: h
    end function
    integer(c_int) function roms_rho_eos_c(h, tidx) bind(C, name='roms_rho_eos')
      import; type(c_ptr), value :: h; integer(c_int), value :: tidx
    end function
    integer(c_int) function roms_prsgrd_c(h) bind(C, name='roms_prsgrd')
      import; type(c_ptr), value :: h
    end function
    integer(c_int) function roms_pre_step3d_c(h, tile) bind(C, name='roms_pre_step3d')
      import; type(c_ptr), value :: h; integer(c_int), value :: tile
    end function
    integer(c_int) function roms_step3d_uv1_c(h, tile) bind(C, name='roms_step3d_uv1')
      import; type(c_ptr), value :: h; integer(c_int), value :: tile
    end function
    integer(c_int) function roms_step2d_c(h) bind(C, name='roms_step2d')
      import; type(c_ptr), value :: h
    end function
    integer(c_int) function roms_step3d_uv2_c(h, tile) bind(C, name='roms_step3d_uv2')
      import; type(c_ptr), value :: h; integer(c_int), value :: tile
    end function
    integer(c_int) function roms_step3d_t_c(h, tile) bind(C, name='roms_step3d_t')
      import; type(c_ptr), value :: h; integer(c_int), value :: tile
    end function
    integer(c_int) function roms_visc3d_c(h) bind(C, name='roms_visc3d')
      import; type(c_ptr), value :: h
    end function
    integer(c_int) function roms_t3dmix_c(h) bind(C, name='roms_t3dmix')
      import; type(c_ptr), value :: h
    end function
    integer(c_int) function roms_lmd_vmix_c(h, tind) bind(C, name='roms_lmd_vmix')
      import; type(c_ptr), value :: h; integer(c_int), value :: tind
    end function
    integer(c_int) function roms_diag_c(h, out4) bind(C, name='roms_diag')
      import; type(c_ptr), value :: h; real(c_double), intent(out) :: out4(4)
    end function
    ! ---- switches and sources the host hands over (set_forces.F:147-152, tides.F, cppdefs LMD_*)
    integer(c_int) function roms_set_kpp(h, on) bind(C, name='roms_set_kpp')
      import; type(c_ptr), value :: h; integer(c_int), value :: on
    end function
    integer(c_int) function roms_set_kpp_opts(h, lmd_convec, Akv_bak, Akt_bak_T, Akt_bak_S) bind(C, name='roms_set_kpp_opts')
      import; type(c_ptr), value :: h; integer(c_int), value :: lmd_convec; real(c_double), value :: Akv_bak, Akt_bak_T, Akt_bak_S
    end function
    integer(c_int) function roms_set_pot_tides(h, on) bind(C, name='roms_set_pot_tides')
      import; type(c_ptr), value :: h; integer(c_int), value :: on
    end function
    integer(c_int) function roms_set_river(h, iriver, vol, trc) bind(C, name='roms_set_river')
      import; type(c_ptr), value :: h; integer(c_int), value :: iriver; real(c_double), value :: vol; real(c_double), intent(in) :: trc(*)
    end function
    integer(c_int) function roms_set_pipe(h, ipipe, prf, trc) bind(C, name='roms_set_pipe')
      import; type(c_ptr), value :: h; integer(c_int), value :: ipipe; real(c_double), intent(in) :: prf(*), trc(*)
    end function
    integer(c_int) function roms_grid_stats(h, out9) bind(C, name='roms_grid_stats')
      import; type(c_ptr), value :: h; real(c_double), intent(out) :: out9(9)
    end function
    ! ---- multi-GPU: mpi_setup keeps computing inode,jnode; rank 0's id is broadcast with MPI_Bcast
    integer(c_int) function roms_comm_unique_id(out, len) bind(C, name='roms_comm_unique_id')
      import; character(c_char), intent(out) :: out(*); integer(c_int), value :: len
    end function
    integer(c_int) function roms_comm_attach(h, id, len, rank, world) bind(C, name='roms_comm_attach')
      import; type(c_ptr), value :: h; character(c_char), intent(in) :: id(*); integer(c_int), value :: len, rank, world
    end function
  end interface

contains

  ! turns a non-zero return code into the reference's collective abort (error_handling_mod.F90:171-374)
  subroutine roms_b200_check(ierr, where)
    use error_handling_mod, only: error_log
    integer(c_int), intent(in) :: ierr
    character(len=*), intent(in) :: where
    if (ierr /= 0) then
      call error_log%raise_global(info='libromsb200: '//where//' failed', context=where)
      call error_log%abort_check()
    end if
  end subroutine roms_b200_check

  ! every shim first tells the library where the driver's index rotation stands (main.F:377-378,425,458-460)
  subroutine roms_b200_sync_indices()
    use scalars, only: iic, iif, nstp, nrhs, nnew, kstp, knew
    integer(c_int) :: ierr
    ierr = roms_set_time_indices(roms_h, int(iic, c_int), int(iif, c_int), int(nstp, c_int), int(nrhs, c_int), &
                                 int(nnew, c_int), int(kstp, c_int), int(knew, c_int))
  end subroutine roms_b200_sync_indices

end module roms_b200_mod
