! delaunay_voronoi_gpu.F90 -- drop-in replacement of src/delaunay_voronoi_mod.F90 whose
! procedures run on the B200 through libfishman_b200.so (include/fishman_b200.h).
!
! What stays exactly as the reference has it ("DV" = /root/reference/src/delaunay_voronoi_mod.F90):
!   * the module name and the public names of DV:15-30: the four procedures, the six derived
!     types and the typed accessors get_DVT / get_DT / get_DE / get_VVT / get_VC / get_VE;
!   * the argument lists of the four procedures (DV:188-195, :226, :358-360, :587-590) and the
!     plug-in contract output_interface (DV:169-181): the output procedure (mpas_mesh_output,
!     src/mpas_mesh_mod.F90:75-83) receives the same six containers of class(*) pointers;
!   * the components of the six types that a plug-in reads (SURVEY.md 8b): ids, coordinates,
!     rings and cross links, in the orders the reference reaches (ring pairing DV:723-738,
!     DE / VE creation order DV:415-459, the ordered VE-in-VC, VE-in-VVT and VC-in-VVT lists the
!     reference arrives at after its fix-ups DV:462-582);
!   * errors end in log_error -> stop 1 (lib/flogger/src/log_mod.F90:99-123).
! What is gone: the serial algorithm (buckets, flips, virtual vertices) and its
! triangulation-only components (cntDT1/2, stub items, subDT, incDVT).  The containers are
! RE-MATERIALISED from the flat arrays the C ABI returns (SURVEY.md 8f row 3).
! New: scvt_lloyd (the reference has no Lloyd loop, SURVEY.md F1) and delaunay_voronoi_final.
!
! Build (GNU): -cpp -O2 -ffree-line-length-none, with lib/container, lib/string, lib/flogger,
! src/const_mod.F90, src/sphere_geometry_mod.F90 of the reference and fishman_gpu_c.F90, linked
! with -lfishman_b200 -lcudart.  NOT COMPILED IN THIS REPOSITORY'S IMAGE NOR ON ITS GPU BOX: neither
! has a Fortran compiler (profiles/r02_probe_fortran_gpubox.txt).  tests/test_fortran_binding.py
! checks mechanically that every fm_* call below exists in the header with that many arguments
! and that the public procedures keep the reference's argument lists.
module delaunay_voronoi_mod

  use, intrinsic :: iso_c_binding
  use flogger
  use container
  use const_mod
  use sphere_geometry_mod
  use fishman_gpu_c

  implicit none

  private

  public delaunay_voronoi_init
  public delaunay_triangulation
  public voronoi_diagram
  public delaunay_voronoi_output
  public delaunay_voronoi_final
  public scvt_lloyd
  public delaunay_vertex_type
  public delaunay_triangle_type
  public delaunay_edge_type
  public voronoi_vertex_type
  public voronoi_cell_type
  public voronoi_edge_type
  public get_DVT
  public get_DT
  public get_DE
  public get_VVT
  public get_VC
  public get_VE

  type, extends(point_type) :: delaunay_vertex_type
    integer :: id = -1
    type(linked_list_type) incDT   ! incident Delaunay triangles, cyclic, counter-clockwise
    type(linked_list_type) linkDVT ! linked Delaunay vertices: linkDVT(j) on the edge incDT(j) | incDT(j+1)
  contains
    procedure :: init => delaunay_vertex_init
  end type delaunay_vertex_type

  type delaunay_triangle_type
    integer :: id = -1
    type(array_type) DVT                         ! three vertices, counter-clockwise seen from outside
    type(array_type) DE
    type(array_type) adjDT                       ! adjDT(k) across the edge opposite DVT(k)
    type(voronoi_vertex_type), pointer :: center => null()
    real(8) :: radius = 0.0d0
  contains
    procedure :: init => delaunay_triangle_init
  end type delaunay_triangle_type

  type delaunay_edge_type
    integer :: id = -1
    type(delaunay_vertex_type), pointer :: DVT1 => null()
    type(delaunay_vertex_type), pointer :: DVT2 => null()
    type(delaunay_triangle_type), pointer :: DT1 => null()
    type(delaunay_triangle_type), pointer :: DT2 => null()
  contains
    procedure :: init => delaunay_edge_init
  end type delaunay_edge_type

  type, extends(point_type) :: voronoi_vertex_type
    integer :: id = -1
    type(delaunay_triangle_type), pointer :: DT => null()
    type(array_type) VE
    type(array_type) VC
  contains
    procedure :: init => voronoi_vertex_init
  end type voronoi_vertex_type

  type voronoi_edge_type
    integer :: id = -1
    type(voronoi_vertex_type), pointer :: VVT1 => null()
    type(voronoi_vertex_type), pointer :: VVT2 => null()
    type(voronoi_cell_type), pointer :: VC1 => null()
    type(voronoi_cell_type), pointer :: VC2 => null()
  contains
    procedure :: init => voronoi_edge_init
  end type voronoi_edge_type

  type voronoi_cell_type
    integer :: id = -1
    type(delaunay_vertex_type), pointer :: center => null()
    type(array_type) VVT
    type(array_type) VE
    type(array_type) adjVC
  contains
    procedure :: init => voronoi_cell_init
  end type voronoi_cell_type

  ! The module state of the reference (DV:107-115), now a cache of what lives on the device
  type(array_type) global_DVT_array
  type(linked_list_type) global_DT_list
  type(array_type) global_DE_array
  type(array_type) global_VVT_array
  type(array_type) global_VE_array
  type(array_type) global_VC_array
  type(array_type) DT_by_id                 ! id -> pointer into global_DT_list

  type(c_ptr), save :: gpu = c_null_ptr     ! the one mesh of the module (the reference is single-shot, SURVEY F5)

  interface get_DVT
    module procedure get_DVT_from_array
    module procedure get_DVT_from_value
  end interface get_DVT
  interface get_DT
    module procedure get_DT_from_array
    module procedure get_DT_from_value
  end interface get_DT
  interface get_DE
    module procedure get_DE_from_array
  end interface get_DE
  interface get_VVT
    module procedure get_VVT_from_array
    module procedure get_VVT_from_value
  end interface get_VVT
  interface get_VC
    module procedure get_VC_from_array
  end interface get_VC
  interface get_VE
    module procedure get_VE_from_array
    module procedure get_VE_from_value
  end interface get_VE

  interface
    subroutine output_interface(DVT_array, DT_list, DE_array, VVT_array, VC_array, VE_array, tag)
      import array_type
      import linked_list_type
      type(array_type), intent(in) :: DVT_array
      type(linked_list_type), intent(in) :: DT_list
      type(array_type), intent(in) :: DE_array
      type(array_type), intent(in) :: VVT_array
      type(array_type), intent(in) :: VC_array
      type(array_type), intent(in) :: VE_array
      character(*), intent(in), optional :: tag
    end subroutine output_interface
  end interface

contains

  subroutine check(rc, where)
    integer(c_int), intent(in) :: rc
    character(*), intent(in) :: where
    if (rc /= 0) call log_error(where // ': ' // fm_error_message(gpu))
  end subroutine check

  ! ------------------------------------------------------------------ delaunay_voronoi_init
  ! Argument list of DV:188-195.  The vertex and cell objects are built as in the reference
  ! (DV:199-220: the plug-in reads VC%center%lon/lat/x/y/z); the coordinates also go to the device.
  ! x, y, z may be non-contiguous sections (gen_mesh passes x(:,1), src/gen_mesh.F90:25): they
  ! are packed into contiguous real(c_double) temporaries before crossing the boundary.
  subroutine delaunay_voronoi_init(num_point, lon, lat, x, y, z)

    integer, intent(in) :: num_point
    real(8), intent(in), optional :: lon(:)
    real(8), intent(in), optional :: lat(:)
    real(8), intent(in), optional :: x(:)
    real(8), intent(in), optional :: y(:)
    real(8), intent(in), optional :: z(:)

    type(delaunay_vertex_type) DVT
    type(voronoi_cell_type) VC
    type(voronoi_cell_type), pointer :: pVC
    real(c_double), allocatable :: a(:), b(:), c(:)
    integer i

    call delaunay_voronoi_final()
    global_DVT_array = array(num_point)
    global_VC_array = array(num_point)
    if (present(lon) .and. present(lat)) then
      allocate(a(num_point), b(num_point))
      a = lon(1:num_point); b = lat(1:num_point)
      if (fm_create_lonlat(int(num_point, c_int32_t), a, b, gpu) /= 0) &
        call log_error('delaunay_voronoi_init: ' // fm_error_message(c_null_ptr))
    else if (present(x) .and. present(y) .and. present(z)) then
      allocate(a(num_point), b(num_point), c(num_point))
      a = x(1:num_point); b = y(1:num_point); c = z(1:num_point)
      if (fm_create(int(num_point, c_int32_t), a, b, c, gpu) /= 0) &
        call log_error('delaunay_voronoi_init: ' // fm_error_message(c_null_ptr))   ! e.g. "DVT i coincides with another DVT", DV:286
    else
      call log_error('delaunay_voronoi_init: need lon, lat or x, y, z')
    end if
    do i = 1, num_point
      if (present(lon) .and. present(lat)) then
        call DVT%init(id=i, lon=lon(i), lat=lat(i))
      else
        call DVT%init(id=i, x=x(i), y=y(i), z=z(i))
      end if
      call global_DVT_array%append(DVT)
      call VC%init(id=i)
      call global_VC_array%append(VC)
      pVC => get_VC(global_VC_array, i)
      pVC%center => get_DVT(global_DVT_array, i)
    end do

  end subroutine delaunay_voronoi_init

  ! Releases the device mesh and the host containers (the reference never frees anything; a
  ! second delaunay_voronoi_init in one process is possible here).
  subroutine delaunay_voronoi_final()

    if (c_associated(gpu)) call fm_destroy(gpu)
    gpu = c_null_ptr
    call clear_topology()
    call global_DVT_array%clear()
    call global_VC_array%clear()

  end subroutine delaunay_voronoi_final

  subroutine clear_topology()

    type(delaunay_vertex_type), pointer :: DVT
    type(voronoi_cell_type), pointer :: VC
    integer i

    call global_DT_list%clear()
    call DT_by_id%clear()
    call global_DE_array%clear()
    call global_VVT_array%clear()
    call global_VE_array%clear()
    do i = 1, global_DVT_array%size
      DVT => get_DVT(global_DVT_array, i)
      call DVT%incDT%clear()
      call DVT%linkDVT%clear()
    end do
    do i = 1, global_VC_array%size
      VC => get_VC(global_VC_array, i)
      call VC%VVT%clear(); call VC%VE%clear(); call VC%adjVC%clear()
      call VC%init(id=i)
      VC%center => get_DVT(global_DVT_array, i)
    end do

  end subroutine clear_topology

  ! ------------------------------------------------------------------ delaunay_triangulation
  ! DV:226-356.  The triangulation is computed on the device (fm_triangulate) and the host
  ! containers are filled from the flat arrays: global_DT_list with triangle ids = list positions
  ! (DV:346-352), DT%DVT counter-clockwise, DT%adjDT(k) opposite DT%DVT(k) (DV:656-669), and for
  ! every vertex the cyclic rings incDT / linkDVT paired as DV:723-738.
  subroutine delaunay_triangulation()

    call check(fm_triangulate(gpu), 'delaunay_triangulation')
    call materialise_triangulation()

  end subroutine delaunay_triangulation

  subroutine materialise_triangulation()

    integer(c_int32_t) nT, nE
    integer(c_int32_t), allocatable :: tri(:,:), adj(:,:), off(:), rtri(:), rnbr(:)
    type(delaunay_triangle_type) DT
    type(delaunay_triangle_type), pointer :: pDT
    type(delaunay_vertex_type), pointer :: DVT
    integer t, k, i, j

    call clear_topology()
    call check(fm_voronoi(gpu, 0_c_int32_t), 'delaunay_triangulation (rings)')
    call check(fm_counts(gpu, nT, nE), 'fm_counts')
    allocate(tri(3,nT), adj(3,nT), off(global_DVT_array%size + 1), rtri(3 * nT), rnbr(3 * nT))
    call check(fm_get_triangles(gpu, tri, adj), 'fm_get_triangles')
    call check(fm_get_rings(gpu, off, rtri, rnbr), 'fm_get_rings')
    DT_by_id = array(int(nT))
    do t = 1, nT
      call DT%init(id=t)
      call global_DT_list%append(DT)
      call DT_by_id%append_ptr(global_DT_list%last_value())
    end do
    do t = 1, nT
      pDT => get_DT(DT_by_id, t)
      do k = 1, 3
        call pDT%DVT%append_ptr(get_DVT(global_DVT_array, int(tri(k,t))))
        call pDT%adjDT%append_ptr(get_DT(DT_by_id, int(adj(k,t))))
      end do
    end do
    do i = 1, global_DVT_array%size
      DVT => get_DVT(global_DVT_array, i)
      do j = off(i) + 1, off(i + 1)
        call DVT%incDT%append_ptr(get_DT(DT_by_id, int(rtri(j))))
        call DVT%linkDVT%append_ptr(get_DVT(global_DVT_array, int(rnbr(j))))
      end do
      call DVT%incDT%cycle()
      call DVT%linkDVT%cycle()
    end do
    call log_notice('Delaunay triangulation is done on the GPU.')

  end subroutine materialise_triangulation

  ! ------------------------------------------------------------------ voronoi_diagram
  ! DV:358-585.  Circumcentres come from the device (they replace calc_circumcircle, DV:1323-1348);
  ! cell rings as DV:394-409.  With all = .true. the Delaunay / Voronoi edges are created in the
  ! order of DV:413-460 (edge k of either kind is the dual of edge k of the other) and the three
  ! ordered lists the reference reaches after its fix-ups DV:462-582 are written directly from the
  ! device tables: VC%VE = edgesOnCell, VVT%VE = edgesOnVertex, VVT%VC = cellsOnVertex (fm_mpas).
  ! No degree limit: the reference's VE_idx(6) (DV:374) does not exist here.
  subroutine voronoi_diagram(all)

    logical, intent(in), optional :: all ! If extract all topological information or not.

    integer(c_int32_t) nT, nE, me
    real(c_double), allocatable :: cc(:,:), ll(:,:), area_cell(:), kite_ring(:), area_tri(:), kite(:,:)
    integer(c_int32_t), allocatable :: off(:), rtri(:), rnbr(:), de(:,:), vv(:,:), vc(:,:), redge(:), cov(:,:), eov(:,:)
    type(voronoi_vertex_type) VVT
    type(voronoi_vertex_type), pointer :: pVVT
    type(voronoi_cell_type), pointer :: VC
    type(delaunay_edge_type) DE
    type(delaunay_edge_type), pointer :: pDE
    type(voronoi_edge_type) VE
    type(voronoi_edge_type), pointer :: pVE
    type(delaunay_vertex_type), pointer :: DVT
    integer i, j, k, t, n

    n = global_DVT_array%size
    call check(fm_counts(gpu, nT, nE), 'fm_counts')
    allocate(cc(3,nT), ll(2,nT), off(n + 1), rtri(3 * nT), rnbr(3 * nT))
    if (present(all)) then
      if (all) call check(fm_voronoi(gpu, 1_c_int32_t), 'voronoi_diagram')
    end if
    call check(fm_get_circumcentres(gpu, cc, ll), 'fm_get_circumcentres')
    call check(fm_get_rings(gpu, off, rtri, rnbr), 'fm_get_rings')
    global_VVT_array = array(int(nT))
    do t = 1, nT
      call VVT%init(id=t)
      VVT%x = cc(1,t); VVT%y = cc(2,t); VVT%z = cc(3,t)
      VVT%lon = ll(1,t); VVT%lat = ll(2,t)
      call global_VVT_array%append(VVT)
      pVVT => get_VVT(global_VVT_array, t)
      pVVT%DT => get_DT(DT_by_id, t)
      pVVT%DT%center => pVVT
    end do
    do i = 1, n
      VC => get_VC(global_VC_array, i)
      do j = off(i) + 1, off(i + 1)
        call VC%adjVC%append_ptr(get_VC(global_VC_array, int(rnbr(j))))
        call VC%VVT%append_ptr(get_VVT(global_VVT_array, int(rtri(j))))
      end do
    end do
    if (.not. present(all)) return
    if (.not. all) return

    allocate(de(2,nE), vv(2,nE), vc(2,nE), redge(3 * nT), area_cell(n), kite_ring(3 * nT))
    allocate(cov(3,nT), eov(3,nT), area_tri(nT), kite(3,nT))
    call check(fm_get_edges(gpu, de, vv, vc), 'fm_get_edges')
    call check(fm_mpas(gpu, 0_c_int32_t, me), 'fm_mpas')
    call check(fm_get_mpas_cells(gpu, redge, area_cell, kite_ring), 'fm_get_mpas_cells')
    call check(fm_get_mpas_vertices(gpu, cov, eov, area_tri, kite), 'fm_get_mpas_vertices')
    global_DE_array = array(int(nE))
    global_VE_array = array(int(nE))
    do k = 1, nE
      call DE%init(id=k)
      DE%DVT1 => get_DVT(global_DVT_array, int(de(1,k)))
      DE%DVT2 => get_DVT(global_DVT_array, int(de(2,k)))
      DE%DT1 => get_DT(DT_by_id, int(vv(1,k)))      ! the two triangles that share the edge
      DE%DT2 => get_DT(DT_by_id, int(vv(2,k)))
      call global_DE_array%append(DE)
      call VE%init(id=k)
      VE%VVT1 => get_VVT(global_VVT_array, int(vv(1,k)))
      VE%VVT2 => get_VVT(global_VVT_array, int(vv(2,k)))
      VE%VC1 => get_VC(global_VC_array, int(vc(1,k)))
      VE%VC2 => get_VC(global_VC_array, int(vc(2,k)))
      call global_VE_array%append(VE)
    end do
    do i = 1, n                                     ! VE in VC, counter-clockwise, VE(j) from VVT(j) to VVT(j+1)
      VC => get_VC(global_VC_array, i)
      do j = off(i) + 1, off(i + 1)
        call VC%VE%append_ptr(get_VE(global_VE_array, int(redge(j))))
      end do
    end do
    do t = 1, nT                                    ! VE and VC around a Voronoi vertex, counter-clockwise
      pVVT => get_VVT(global_VVT_array, t)
      do k = 1, 3
        call pVVT%VE%append_ptr(get_VE(global_VE_array, int(eov(k,t))))
        call pVVT%VC%append_ptr(get_VC(global_VC_array, int(cov(k,t))))
      end do
    end do
    do t = 1, nT                                    ! DT%DE: the edge opposite DT%DVT(k)
      pVVT => get_VVT(global_VVT_array, t)
      do k = 1, 3
        do j = 1, 3
          pVE => get_VE(pVVT%VE, j)
          pDE => get_DE(global_DE_array, pVE%id)
          DVT => get_DVT(pVVT%DT%DVT, k)
          if (pDE%DVT1%id /= DVT%id .and. pDE%DVT2%id /= DVT%id) call pVVT%DT%DE%append_ptr(pDE)
        end do
      end do
    end do

  end subroutine voronoi_diagram

  ! ------------------------------------------------------------------ delaunay_voronoi_output
  ! DV:587-593, unchanged: the plug-in gets the six containers.
  subroutine delaunay_voronoi_output(output)

    procedure(output_interface) output

    call output(global_DVT_array, global_DT_list, global_DE_array, global_VVT_array, global_VC_array, global_VE_array)

  end subroutine delaunay_voronoi_output

  ! ------------------------------------------------------------------ scvt_lloyd (new)
  ! The Lloyd loop the reference lacks (SURVEY.md 7.3): at most max_iter iterations of
  ! triangulation -> circumcentres -> density-weighted centroids -> generator update on the
  ! device, stopping when the largest chord displacement is <= tol.  Afterwards the vertex
  ! objects carry the relaxed coordinates and the containers hold the triangulation of the final
  ! generators; call voronoi_diagram(all) and delaunay_voronoi_output as usual.
  subroutine scvt_lloyd(max_iter, tol, iters, max_disp, density)

    integer, intent(in) :: max_iter
    real(8), intent(in) :: tol
    integer, intent(out) :: iters
    real(8), intent(out) :: max_disp
    type(fm_density), intent(in), optional :: density

    type(fm_density) rho                     ! default initialisation = uniform density
    integer(c_int32_t) it
    real(c_double) md
    real(c_double), allocatable :: a(:), b(:), c(:)
    type(delaunay_vertex_type), pointer :: DVT
    integer i, n

    if (present(density)) rho = density
    call check(fm_lloyd(gpu, rho, int(max_iter, c_int32_t), real(tol, c_double), 1_c_int32_t, it, md), 'scvt_lloyd')
    iters = it
    max_disp = md
    n = global_DVT_array%size
    allocate(a(n), b(n), c(n))
    call check(fm_get_generators(gpu, a, b, c), 'fm_get_generators')
    do i = 1, n
      DVT => get_DVT(global_DVT_array, i)
      call DVT%init(id=i, x=a(i), y=b(i), z=c(i))
    end do
    call check(fm_retriangulate(gpu, 1_c_int32_t), 'scvt_lloyd (final triangulation)')
    call materialise_triangulation()

  end subroutine scvt_lloyd

  ! ------------------------------------------------------------------ type-bound initialisers
  subroutine delaunay_vertex_init(this, id, lon, lat, x, y, z)

    class(delaunay_vertex_type), intent(inout) :: this
    integer, intent(in) :: id
    real(8), intent(in), optional :: lon
    real(8), intent(in), optional :: lat
    real(8), intent(in), optional :: x
    real(8), intent(in), optional :: y
    real(8), intent(in), optional :: z

    this%id = id
    if (present(lon) .and. present(lat)) then
      this%lon = lon; this%lat = lat
      call cartesian_transform(this)           ! SG:87-98
    else if (present(x) .and. present(y) .and. present(z)) then
      this%x = x; this%y = y; this%z = z
      call inverse_cartesian_transform(this)   ! SG:100-121: lon in [0, 2 pi)
    end if

  end subroutine delaunay_vertex_init

  subroutine delaunay_triangle_init(this, id)

    class(delaunay_triangle_type), intent(inout) :: this
    integer, intent(in), optional :: id

    if (present(id)) this%id = id
    this%DVT = array(3)
    this%DE = array(3)
    this%adjDT = array(3)
    nullify(this%center)

  end subroutine delaunay_triangle_init

  subroutine delaunay_edge_init(this, id)

    class(delaunay_edge_type), intent(inout) :: this
    integer, intent(in) :: id

    this%id = id
    nullify(this%DVT1, this%DVT2, this%DT1, this%DT2)

  end subroutine delaunay_edge_init

  subroutine voronoi_vertex_init(this, id)

    class(voronoi_vertex_type), intent(inout) :: this
    integer, intent(in) :: id

    this%id = id
    this%VE = array(3)
    this%VC = array(3)
    nullify(this%DT)

  end subroutine voronoi_vertex_init

  subroutine voronoi_edge_init(this, id)

    class(voronoi_edge_type), intent(inout) :: this
    integer, intent(in) :: id

    this%id = id
    nullify(this%VVT1, this%VVT2, this%VC1, this%VC2)

  end subroutine voronoi_edge_init

  subroutine voronoi_cell_init(this, id)

    class(voronoi_cell_type), intent(inout) :: this
    integer, intent(in) :: id

    this%id = id
    this%VVT = array(6)      ! grows on append: no degree limit here
    this%VE = array(6)
    this%adjVC = array(6)

  end subroutine voronoi_cell_init

  ! ------------------------------------------------------------------ typed accessors (DV:15-30)
  function get_DVT_from_value(value) result(res)
    class(*), intent(in), target :: value
    type(delaunay_vertex_type), pointer :: res
    res => null()
    select type (value)
    type is (delaunay_vertex_type)
      res => value
    end select
  end function get_DVT_from_value

  function get_DVT_from_array(array, index) result(res)
    type(array_type), intent(in) :: array
    integer, intent(in) :: index
    type(delaunay_vertex_type), pointer :: res
    res => get_DVT_from_value(array%value_at(index))
  end function get_DVT_from_array

  function get_DT_from_value(value) result(res)
    class(*), intent(in), target :: value
    type(delaunay_triangle_type), pointer :: res
    res => null()
    select type (value)
    type is (delaunay_triangle_type)
      res => value
    end select
  end function get_DT_from_value

  function get_DT_from_array(array, index) result(res)
    type(array_type), intent(in) :: array
    integer, intent(in) :: index
    type(delaunay_triangle_type), pointer :: res
    res => get_DT_from_value(array%value_at(index))
  end function get_DT_from_array

  function get_DE_from_array(array, index) result(res)
    type(array_type), intent(in) :: array
    integer, intent(in) :: index
    type(delaunay_edge_type), pointer :: res
    res => null()
    select type (val => array%value_at(index))
    type is (delaunay_edge_type)
      res => val
    end select
  end function get_DE_from_array

  function get_VVT_from_value(value) result(res)
    class(*), intent(in), target :: value
    type(voronoi_vertex_type), pointer :: res
    res => null()
    select type (value)
    type is (voronoi_vertex_type)
      res => value
    end select
  end function get_VVT_from_value

  function get_VVT_from_array(array, index) result(res)
    type(array_type), intent(in) :: array
    integer, intent(in) :: index
    type(voronoi_vertex_type), pointer :: res
    res => get_VVT_from_value(array%value_at(index))
  end function get_VVT_from_array

  function get_VC_from_array(array, index) result(res)
    type(array_type), intent(in) :: array
    integer, intent(in) :: index
    type(voronoi_cell_type), pointer :: res
    res => null()
    select type (val => array%value_at(index))
    type is (voronoi_cell_type)
      res => val
    end select
  end function get_VC_from_array

  function get_VE_from_value(value) result(res)
    class(*), intent(in), target :: value
    type(voronoi_edge_type), pointer :: res
    res => null()
    select type (value)
    type is (voronoi_edge_type)
      res => value
    end select
  end function get_VE_from_value

  function get_VE_from_array(array, index) result(res)
    type(array_type), intent(in) :: array
    integer, intent(in) :: index
    type(voronoi_edge_type), pointer :: res
    res => get_VE_from_value(array%value_at(index))
  end function get_VE_from_array

end module delaunay_voronoi_mod
