! roms_b200_shims.F90 -- the reference's tile subroutines (same names, same arguments) as thin shims on libromsb200.so.
! Each replaces the body of the subroutine cited beside it; main.F:397-479 calls them unchanged.

subroutine step2d                                   ! step2d_FB.F:3
  use roms_b200_mod
  call roms_b200_sync_indices()
  call roms_b200_check(roms_step2d_c(roms_h), 'step2d')
end subroutine step2d

subroutine pre_step3d(tile)                         ! pre_step3d4S.F:4
  use roms_b200_mod
  integer, intent(in) :: tile
  call roms_b200_sync_indices()
  call roms_b200_check(roms_pre_step3d_c(roms_h, int(tile, c_int)), 'pre_step3d')
end subroutine pre_step3d

subroutine prsgrd                                   ! prsgrd.F:4
  use roms_b200_mod
  call roms_b200_sync_indices()
  call roms_b200_check(roms_prsgrd_c(roms_h), 'prsgrd')
end subroutine prsgrd

subroutine rho_eos(tidx)                            ! rho_eos.F:6
  use roms_b200_mod
  integer, intent(in) :: tidx
  call roms_b200_sync_indices()
  call roms_b200_check(roms_rho_eos_c(roms_h, int(tidx, c_int)), 'rho_eos')
end subroutine rho_eos

subroutine step3d_uv1(tile)                         ! step3d_uv1.F:5
  use roms_b200_mod
  integer, intent(in) :: tile
  call roms_b200_sync_indices()
  call roms_b200_check(roms_step3d_uv1_c(roms_h, int(tile, c_int)), 'step3d_uv1')
end subroutine step3d_uv1

subroutine step3d_uv2(tile)                         ! step3d_uv2.F:6
  use roms_b200_mod
  integer, intent(in) :: tile
  call roms_b200_sync_indices()
  call roms_b200_check(roms_step3d_uv2_c(roms_h, int(tile, c_int)), 'step3d_uv2')
end subroutine step3d_uv2

subroutine step3d_t(tile)                           ! step3d_t_ISO.F:20
  use roms_b200_mod
  integer, intent(in) :: tile
  call roms_b200_sync_indices()
  call roms_b200_check(roms_step3d_t_c(roms_h, int(tile, c_int)), 'step3d_t')
end subroutine step3d_t

subroutine omega                                    ! omega.F:4
  use roms_b200_mod
  call roms_b200_sync_indices()
  call roms_b200_check(roms_omega_c(roms_h), 'omega')
end subroutine omega

subroutine set_depth(tile)                          ! set_depth.F:4
  use roms_b200_mod
  integer, intent(in) :: tile
  call roms_b200_sync_indices()
  call roms_b200_check(roms_set_depth_c(roms_h, int(tile, c_int)), 'set_depth')
end subroutine set_depth

subroutine set_HUV                                  ! set_depth.F:190
  use roms_b200_mod
  call roms_b200_sync_indices()
  call roms_b200_check(roms_set_huv_c(roms_h), 'set_HUV')
end subroutine set_HUV

subroutine set_HUV1(tile)                           ! set_depth.F:239
  use roms_b200_mod
  integer, intent(in) :: tile
  call roms_b200_sync_indices()
  call roms_b200_check(roms_set_huv1_c(roms_h, int(tile, c_int)), 'set_HUV1')
end subroutine set_HUV1

! with UV_VIS2 / TS_DIF2 / LMD_MIXING also (main.F:408-410,434-436,448-450,471-473):
subroutine visc3d                                   ! visc3d_S.F:4
  use roms_b200_mod
  call roms_b200_sync_indices()
  call roms_b200_check(roms_visc3d_c(roms_h), 'visc3d')
end subroutine visc3d

subroutine t3dmix                                   ! t3dmix_S.F:4
  use roms_b200_mod
  call roms_b200_sync_indices()
  call roms_b200_check(roms_t3dmix_c(roms_h), 't3dmix')
end subroutine t3dmix

subroutine lmd_vmix(tind)                           ! lmd_vmix.F:4
  use roms_b200_mod
  integer, intent(in) :: tind
  call roms_b200_sync_indices()
  call roms_b200_check(roms_lmd_vmix_c(roms_h, int(tind, c_int)), 'lmd_vmix')
end subroutine lmd_vmix
