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

! Open boundaries (cfg%obc_* = 1): hand the lines that set_bry_all has just filled (boundary.F:81-338) to the library.  Call it right
! after each `call set_bry_all` of roms_step (main.F:392, :440); the boundary conditions themselves (zetabc, u2dbc, v2dbc, u3dbc, v3dbc,
! t3dbc) run on the device inside step2d, pre_step3d, step3d_uv2 and step3d_t.  The arrays keep the reference's extents
! (zeta_west(0:Mm+1), u_west(0:Mm+1,N), t_west(0:Mm+1,N,NT), ... _south(0:Lm+1,...)).
subroutine roms_b200_push_bry
  use roms_b200_mod
  use param, only: nt
  use boundary, only: zeta_west, ubar_west, vbar_west, u_west, v_west, t_west,     &
                      zeta_east, ubar_east, vbar_east, u_east, v_east, t_east,     &
                      zeta_south, ubar_south, vbar_south, u_south, v_south, t_south, &
                      zeta_north, ubar_north, vbar_north, u_north, v_north, t_north
  integer :: itrc
#ifdef OBC_WEST
  call roms_b200_check(roms_push_async(roms_h, ROMS_ZETA_WEST, 1, 1, c_loc(zeta_west(0))), 'push zeta_west')
  call roms_b200_check(roms_push_async(roms_h, ROMS_UBAR_WEST, 1, 1, c_loc(ubar_west(0))), 'push ubar_west')
  call roms_b200_check(roms_push_async(roms_h, ROMS_VBAR_WEST, 1, 1, c_loc(vbar_west(0))), 'push vbar_west')
  call roms_b200_check(roms_push_async(roms_h, ROMS_U_WEST, 1, 1, c_loc(u_west(0,1))), 'push u_west')
  call roms_b200_check(roms_push_async(roms_h, ROMS_V_WEST, 1, 1, c_loc(v_west(0,1))), 'push v_west')
  do itrc = 1, nt
    call roms_b200_check(roms_push_async(roms_h, ROMS_T_WEST, 1, int(itrc, c_int), c_loc(t_west(0,1,itrc))), 'push t_west')
  enddo
#endif
#ifdef OBC_EAST
  call roms_b200_check(roms_push_async(roms_h, ROMS_ZETA_EAST, 1, 1, c_loc(zeta_east(0))), 'push zeta_east')
  call roms_b200_check(roms_push_async(roms_h, ROMS_UBAR_EAST, 1, 1, c_loc(ubar_east(0))), 'push ubar_east')
  call roms_b200_check(roms_push_async(roms_h, ROMS_VBAR_EAST, 1, 1, c_loc(vbar_east(0))), 'push vbar_east')
  call roms_b200_check(roms_push_async(roms_h, ROMS_U_EAST, 1, 1, c_loc(u_east(0,1))), 'push u_east')
  call roms_b200_check(roms_push_async(roms_h, ROMS_V_EAST, 1, 1, c_loc(v_east(0,1))), 'push v_east')
  do itrc = 1, nt
    call roms_b200_check(roms_push_async(roms_h, ROMS_T_EAST, 1, int(itrc, c_int), c_loc(t_east(0,1,itrc))), 'push t_east')
  enddo
#endif
#ifdef OBC_SOUTH
  call roms_b200_check(roms_push_async(roms_h, ROMS_ZETA_SOUTH, 1, 1, c_loc(zeta_south(0))), 'push zeta_south')
  call roms_b200_check(roms_push_async(roms_h, ROMS_UBAR_SOUTH, 1, 1, c_loc(ubar_south(0))), 'push ubar_south')
  call roms_b200_check(roms_push_async(roms_h, ROMS_VBAR_SOUTH, 1, 1, c_loc(vbar_south(0))), 'push vbar_south')
  call roms_b200_check(roms_push_async(roms_h, ROMS_U_SOUTH, 1, 1, c_loc(u_south(0,1))), 'push u_south')
  call roms_b200_check(roms_push_async(roms_h, ROMS_V_SOUTH, 1, 1, c_loc(v_south(0,1))), 'push v_south')
  do itrc = 1, nt
    call roms_b200_check(roms_push_async(roms_h, ROMS_T_SOUTH, 1, int(itrc, c_int), c_loc(t_south(0,1,itrc))), 'push t_south')
  enddo
#endif
#ifdef OBC_NORTH
  call roms_b200_check(roms_push_async(roms_h, ROMS_ZETA_NORTH, 1, 1, c_loc(zeta_north(0))), 'push zeta_north')
  call roms_b200_check(roms_push_async(roms_h, ROMS_UBAR_NORTH, 1, 1, c_loc(ubar_north(0))), 'push ubar_north')
  call roms_b200_check(roms_push_async(roms_h, ROMS_VBAR_NORTH, 1, 1, c_loc(vbar_north(0))), 'push vbar_north')
  call roms_b200_check(roms_push_async(roms_h, ROMS_U_NORTH, 1, 1, c_loc(u_north(0,1))), 'push u_north')
  call roms_b200_check(roms_push_async(roms_h, ROMS_V_NORTH, 1, 1, c_loc(v_north(0,1))), 'push v_north')
  do itrc = 1, nt
    call roms_b200_check(roms_push_async(roms_h, ROMS_T_NORTH, 1, int(itrc, c_int), c_loc(t_north(0,1,itrc))), 'push t_north')
  enddo
#endif
end subroutine roms_b200_push_bry
