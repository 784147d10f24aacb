! fishman_gpu_c.F90 -- ISO_C_BINDING view of include/fishman_b200.h.
!
! This is the thin bind(C) layer the north star asks for: the Fortran host
! (gen_mesh, the output_interface plug-in contract, mpas_mesh_mod) stays as it is
! and the bodies of delaunay_voronoi_init / delaunay_triangulation /
! voronoi_diagram in src/delaunay_voronoi_mod.F90 call these entry points instead
! of running the serial algorithm (see INTEGRATION.md for the replacement bodies).
!
! NOT COMPILED IN THIS REPOSITORY'S IMAGE: no Fortran compiler exists there.  The
! signatures below are checked mechanically against the C header by
! tests/test_fortran_binding.py (names, argument counts, value/reference passing).
module fishman_gpu_c

  use, intrinsic :: iso_c_binding

  implicit none

  private
  public :: fm_density
  public :: fm_create, fm_create_lonlat, fm_create_dist, fm_destroy, fm_last_error
  public :: fm_triangulate, fm_retriangulate, fm_counts, fm_get_triangles
  public :: fm_voronoi, fm_get_circumcentres, fm_get_rings, fm_get_edges
  public :: fm_mpas, fm_get_mpas_edges, fm_get_mpas_cells, fm_get_mpas_vertices, fm_get_mpas_edges_on_edge
  public :: fm_lloyd, fm_lloyd_local, fm_generator_block, fm_peer_handle, fm_peer_attach
  public :: fm_get_generators, fm_set_generators, fm_density_eval
  public :: fm_quality_stats, fm_quality, fm_block_ids, fm_get_generators_block, fm_set_generators_block, fm_fp64_peak
  public :: fm_quality_static, fm_star_stats
  public :: fm_error_message

  ! typedef struct fm_density { int32_t kind; double cx, cy, cz, beta, alpha, gamma; }
  type, bind(C) :: fm_density
    integer(c_int32_t) :: kind = 0
    real(c_double) :: cx = 0.0d0, cy = 0.0d0, cz = 1.0d0
    real(c_double) :: beta = 0.0d0, alpha = 1.0d0, gamma = 1.0d0
  end type fm_density

  ! typedef struct fm_quality_stats (mesh-quality statistics, SURVEY.md 7.3)
  type, bind(C) :: fm_quality_stats
    integer(c_int64_t) :: degree_hist(16)
    real(c_double) :: area_cell_min, area_cell_max, area_cell_sum
    real(c_double) :: area_tri_min, area_tri_max, area_tri_sum
    real(c_double) :: angle_min, angle_max
    real(c_double) :: max_disp
  end type fm_quality_stats

  interface

    ! delaunay_voronoi_init(num_point, x=, y=, z=), DV:188-224
    integer(c_int) function fm_create(n, x, y, z, h) bind(C, name='fm_create')
      import :: c_int, c_int32_t, c_double, c_ptr
      integer(c_int32_t), value :: n
      real(c_double), intent(in) :: x(*), y(*), z(*)
      type(c_ptr), intent(out) :: h
    end function fm_create

    ! delaunay_voronoi_init(num_point, lon=, lat=), DV:204-211
    integer(c_int) function fm_create_lonlat(n, lon, lat, h) bind(C, name='fm_create_lonlat')
      import :: c_int, c_int32_t, c_double, c_ptr
      integer(c_int32_t), value :: n
      real(c_double), intent(in) :: lon(*), lat(*)
      type(c_ptr), intent(out) :: h
    end function fm_create_lonlat

    integer(c_int) function fm_create_dist(n, x, y, z, rank, size, h) bind(C, name='fm_create_dist')
      import :: c_int, c_int32_t, c_double, c_ptr
      integer(c_int32_t), value :: n, rank, size
      real(c_double), intent(in) :: x(*), y(*), z(*)
      type(c_ptr), intent(out) :: h
    end function fm_create_dist

    subroutine fm_destroy(h) bind(C, name='fm_destroy')
      import :: c_ptr
      type(c_ptr), value :: h
    end subroutine fm_destroy

    type(c_ptr) function fm_last_error(h) bind(C, name='fm_last_error')
      import :: c_ptr
      type(c_ptr), value :: h
    end function fm_last_error

    ! delaunay_triangulation(), DV:226-356
    integer(c_int) function fm_triangulate(h) bind(C, name='fm_triangulate')
      import :: c_int, c_ptr
      type(c_ptr), value :: h
    end function fm_triangulate

    integer(c_int) function fm_retriangulate(h, warm) bind(C, name='fm_retriangulate')
      import :: c_int, c_int32_t, c_ptr
      type(c_ptr), value :: h
      integer(c_int32_t), value :: warm
    end function fm_retriangulate

    integer(c_int) function fm_counts(h, n_tri, n_edge) bind(C, name='fm_counts')
      import :: c_int, c_int32_t, c_ptr
      type(c_ptr), value :: h
      integer(c_int32_t), intent(out) :: n_tri, n_edge
    end function fm_counts

    ! tri(3,T), adj(3,T): Fortran column-major = the C layout
    integer(c_int) function fm_get_triangles(h, tri, adj) bind(C, name='fm_get_triangles')
      import :: c_int, c_int32_t, c_ptr
      type(c_ptr), value :: h
      integer(c_int32_t), intent(out) :: tri(3, *), adj(3, *)
    end function fm_get_triangles

    ! voronoi_diagram(all), DV:358-585
    integer(c_int) function fm_voronoi(h, all) bind(C, name='fm_voronoi')
      import :: c_int, c_int32_t, c_ptr
      type(c_ptr), value :: h
      integer(c_int32_t), value :: all
    end function fm_voronoi

    integer(c_int) function fm_get_circumcentres(h, xyz, lonlat) bind(C, name='fm_get_circumcentres')
      import :: c_int, c_double, c_ptr
      type(c_ptr), value :: h
      real(c_double), intent(out) :: xyz(3, *), lonlat(2, *)
    end function fm_get_circumcentres

    integer(c_int) function fm_get_rings(h, off, ring_tri, ring_nbr) bind(C, name='fm_get_rings')
      import :: c_int, c_int32_t, c_ptr
      type(c_ptr), value :: h
      integer(c_int32_t), intent(out) :: off(*), ring_tri(*), ring_nbr(*)
    end function fm_get_rings

    integer(c_int) function fm_get_edges(h, de, ve_vvt, ve_vc) bind(C, name='fm_get_edges')
      import :: c_int, c_int32_t, c_ptr
      type(c_ptr), value :: h
      integer(c_int32_t), intent(out) :: de(2, *), ve_vvt(2, *), ve_vc(2, *)
    end function fm_get_edges

    ! Lloyd loop: no reference procedure (SURVEY.md F1)
    ! mpas_mesh_adapt + calc_weightsOnEdge on the device (src/mpas_mesh_mod.F90:181-345, :442-561)
    integer(c_int) function fm_mpas(h, strict, max_edges) bind(C, name='fm_mpas')
      import :: c_int, c_int32_t, c_ptr
      type(c_ptr), value :: h
      integer(c_int32_t), value :: strict
      integer(c_int32_t), intent(out) :: max_edges
    end function fm_mpas

    integer(c_int) function fm_get_mpas_edges(h, xyz, lonlat, dv, dc, dv1, dv2, angle) bind(C, name='fm_get_mpas_edges')
      import :: c_int, c_double, c_ptr
      type(c_ptr), value :: h
      real(c_double), intent(out) :: xyz(3,*), lonlat(2,*), dv(*), dc(*), dv1(*), dv2(*), angle(*)
    end function fm_get_mpas_edges

    integer(c_int) function fm_get_mpas_cells(h, edges_on_cell, area_cell, kite_on_cell) bind(C, name='fm_get_mpas_cells')
      import :: c_int, c_int32_t, c_double, c_ptr
      type(c_ptr), value :: h
      integer(c_int32_t), intent(out) :: edges_on_cell(*)
      real(c_double), intent(out) :: area_cell(*), kite_on_cell(*)
    end function fm_get_mpas_cells

    integer(c_int) function fm_get_mpas_vertices(h, cells_on_vertex, edges_on_vertex, area_triangle, kite) &
        bind(C, name='fm_get_mpas_vertices')
      import :: c_int, c_int32_t, c_double, c_ptr
      type(c_ptr), value :: h
      integer(c_int32_t), intent(out) :: cells_on_vertex(3,*), edges_on_vertex(3,*)
      real(c_double), intent(out) :: area_triangle(*), kite(3,*)
    end function fm_get_mpas_vertices

    integer(c_int) function fm_get_mpas_edges_on_edge(h, n_edges_on_edge, edges_on_edge, weights_on_edge) &
        bind(C, name='fm_get_mpas_edges_on_edge')
      import :: c_int, c_int32_t, c_double, c_ptr
      type(c_ptr), value :: h
      integer(c_int32_t), intent(out) :: n_edges_on_edge(*), edges_on_edge(*)
      real(c_double), intent(out) :: weights_on_edge(*)
    end function fm_get_mpas_edges_on_edge

    integer(c_int) function fm_lloyd(h, rho, max_iter, tol, warm, iters, max_disp) bind(C, name='fm_lloyd')
      import :: c_int, c_int32_t, c_double, c_ptr, fm_density
      type(c_ptr), value :: h
      type(fm_density), intent(in) :: rho
      integer(c_int32_t), value :: max_iter, warm
      real(c_double), value :: tol
      integer(c_int32_t), intent(out) :: iters
      real(c_double), intent(out) :: max_disp
    end function fm_lloyd

    integer(c_int) function fm_lloyd_local(h, rho, warm, max_disp) bind(C, name='fm_lloyd_local')
      import :: c_int, c_int32_t, c_double, c_ptr, fm_density
      type(c_ptr), value :: h
      type(fm_density), intent(in) :: rho
      integer(c_int32_t), value :: warm
      real(c_double), intent(out) :: max_disp
    end function fm_lloyd_local

    ! rho at n unit vectors, evaluated by the device code of the Lloyd kernel (diagnostics)
    integer(c_int) function fm_density_eval(rho, n, x, y, z, out) bind(C, name='fm_density_eval')
      import :: c_int, c_int64_t, c_double, fm_density
      type(fm_density), intent(in) :: rho
      integer(c_int64_t), value :: n
      real(c_double), intent(in) :: x(*), y(*), z(*)
      real(c_double), intent(out) :: out(*)
    end function fm_density_eval

    integer(c_int) function fm_generator_block(h, d_records, n_padded, begin, count) &
        bind(C, name='fm_generator_block')
      import :: c_int, c_int64_t, c_ptr
      type(c_ptr), value :: h
      type(c_ptr), intent(out) :: d_records
      integer(c_int64_t), intent(out) :: n_padded, begin, count
    end function fm_generator_block

    ! 80-byte blob (CUDA IPC handle + device UUID) of this rank; exchange with MPI_Allgather, then attach all
    integer(c_int) function fm_peer_handle(h, out80) bind(C, name='fm_peer_handle')
      import :: c_int, c_char, c_ptr
      type(c_ptr), value :: h
      character(kind=c_char), intent(out) :: out80(80)
    end function fm_peer_handle

    integer(c_int) function fm_peer_attach(h, handles) bind(C, name='fm_peer_attach')
      import :: c_int, c_char, c_ptr
      type(c_ptr), value :: h
      character(kind=c_char), intent(in) :: handles(*)
    end function fm_peer_attach

    integer(c_int) function fm_get_generators(h, x, y, z) bind(C, name='fm_get_generators')
      import :: c_int, c_double, c_ptr
      type(c_ptr), value :: h
      real(c_double), intent(out) :: x(*), y(*), z(*)
    end function fm_get_generators

    integer(c_int) function fm_set_generators(h, x, y, z) bind(C, name='fm_set_generators')
      import :: c_int, c_double, c_ptr
      type(c_ptr), value :: h
      real(c_double), intent(in) :: x(*), y(*), z(*)
    end function fm_set_generators

    integer(c_int) function fm_quality(h, rho, q) bind(C, name='fm_quality')
      import :: c_int, c_ptr, fm_density, fm_quality_stats
      type(c_ptr), value :: h
      type(fm_density), intent(in) :: rho
      type(fm_quality_stats), intent(out) :: q
    end function fm_quality

    ! fm_quality under the uniform density without any device allocation (no counterpart in the reference; scratch in
    ! a static device array of the library, at most 4,194,304 generators, the triangulation must be current)
    integer(c_int) function fm_quality_static(h, q) bind(C, name='fm_quality_static')
      import :: c_int, c_ptr, fm_quality_stats
      type(c_ptr), value :: h
      type(fm_quality_stats), intent(out) :: q
    end function fm_quality_static

    ! how the last cold build was made (by stars or by insertion + flips): ten 64-bit words, include/fishman_b200.h
    integer(c_int) function fm_star_stats(h, out) bind(C, name='fm_star_stats')
      import :: c_int, c_int64_t, c_ptr
      type(c_ptr), value :: h
      integer(c_int64_t), intent(out) :: out(10)
    end function fm_star_stats

    ! one rank's block of a sharded mesh (fm_create_dist): ids, download, upload
    integer(c_int) function fm_block_ids(h, ids) bind(C, name='fm_block_ids')
      import :: c_int, c_int32_t, c_ptr
      type(c_ptr), value :: h
      integer(c_int32_t), intent(out) :: ids(*)
    end function fm_block_ids

    integer(c_int) function fm_get_generators_block(h, x, y, z) bind(C, name='fm_get_generators_block')
      import :: c_int, c_double, c_ptr
      type(c_ptr), value :: h
      real(c_double), intent(out) :: x(*), y(*), z(*)
    end function fm_get_generators_block

    integer(c_int) function fm_set_generators_block(h, x, y, z) bind(C, name='fm_set_generators_block')
      import :: c_int, c_double, c_ptr
      type(c_ptr), value :: h
      real(c_double), intent(in) :: x(*), y(*), z(*)
    end function fm_set_generators_block

    integer(c_int) function fm_fp64_peak(min_ms, tflops) bind(C, name='fm_fp64_peak')
      import :: c_int, c_double
      real(c_double), value :: min_ms
      real(c_double), intent(out) :: tflops
    end function fm_fp64_peak

  end interface

contains

  ! C string of fm_last_error -> Fortran string (for log_error)
  function fm_error_message(h) result(msg)
    type(c_ptr), intent(in) :: h
    character(:), allocatable :: msg
    type(c_ptr) :: p
    character(kind=c_char), pointer :: s(:)
    integer :: n
    p = fm_last_error(h)
    if (.not. c_associated(p)) then
      msg = 'unknown error'
      return
    end if
    call c_f_pointer(p, s, [512])
    n = 0
    do while (n < 512)
      if (s(n + 1) == c_null_char) exit
      n = n + 1
    end do
    allocate(character(n) :: msg)
    msg = transfer(s(1:n), msg)
  end function fm_error_message

end module fishman_gpu_c
