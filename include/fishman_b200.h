/*
 * fishman_b200.h -- C ABI of the B200 (sm_100a) implementation of fishman's SCVT
 * inner loop: spherical Delaunay triangulation, dual Voronoi diagram and the
 * Lloyd centroid update.
 *
 * This is the boundary a Fortran host binds with ISO_C_BINDING (the stub is in
 * fishman_b200/fortran/fishman_gpu_c.F90 and INTEGRATION.md).  Each entry point
 * names the reference procedure it replaces; "DV" is
 * /root/reference/src/delaunay_voronoi_mod.F90, "SG" is
 * /root/reference/src/sphere_geometry_mod.F90, "MM" is
 * /root/reference/src/mpas_mesh_mod.F90.
 *
 * Conventions (those of the reference):
 *   - every id that crosses the boundary is 1-based; cell / generator ids are the
 *     input order (DV:341-344);
 *   - triangles are counter-clockwise seen from outside the sphere (DV:1882-1885);
 *   - adj[k] is the triangle across the edge opposite vertex k (DV:656-669);
 *   - ring_nbr[j] lies on the edge shared by ring_tri[j] and ring_tri[j+1]
 *     (DV:723-738); rings start at their smallest neighbour id;
 *   - lon is in [0, 2 pi) (SG:119);
 *   - T = 2N-4 triangles, E = 3N-6 edges (euler_formula, SG:57-71).
 * Every function returns 0 on success; on failure it returns non-zero and
 * fm_last_error() describes it (the Fortran shim turns that into log_error ->
 * stop 1, the reference's only error convention, lib/flogger/src/log_mod.F90:99).
 * There is no CPU fallback: without a CUDA device fm_create fails.
 * A handle is not thread-safe; distinct handles are independent.
 */
#ifndef FISHMAN_B200_H
#define FISHMAN_B200_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct fm_handle fm_handle;

/* Density function of the Lloyd step (the reference has none).
 * kind 0: rho = 1.  kind 1: rho(x) = (1-gamma)/2 (tanh((beta - d(x,c))/alpha) + 1) + gamma,
 * d = great-circle distance to the unit vector c (calc_arc_length, SG:289-296). */
typedef struct fm_density {
  int32_t kind;
  double cx, cy, cz, beta, alpha, gamma;
} fm_density;

/* rho at n unit vectors, evaluated on the device (host pointers in and out; no mesh needed), for
 * diagnostics and for the parity tests of the density function itself.  By default through the
 * formula (device acos / exp without library calls, relative error ~1e-16); with the environment
 * variable FM_DENSITY_TABLE=1 through the staged table the Lloyd kernel reads (rho as a piecewise
 * degree-7 polynomial of the chord to the centre, checked on the host to 2e-13 per interval;
 * intervals that miss it are evaluated by the formula).  FM_DENSITY_TABLE=0 makes the Lloyd kernel
 * use the formula as well. */
int fm_density_eval(const fm_density* rho, int64_t n, const double* x, const double* y, const double* z,
                    double* out);

/* delaunay_voronoi_init(num_point, x=, y=, z=)  DV:188-224.  Copies the n generators to
 * the device (contiguous arrays: the shim packs strided sections first), orders them
 * along a space-filling curve internally and rejects duplicate points (DV:286-288). */
int fm_create(int32_t n, const double* x, const double* y, const double* z, fm_handle** out);
/* delaunay_voronoi_init(num_point, lon=, lat=)  DV:204-211 (cartesian_transform, SG:87-98) */
int fm_create_lonlat(int32_t n, const double* lon, const double* lat, fm_handle** out);
/* Sharded variant: this process is rank `rank` of `size`; the Lloyd centroid stage then
 * updates only the rank's contiguous block of the internal (space-filling-curve)
 * order, see fm_lloyd_local / fm_generator_block. */
int fm_create_dist(int32_t n, const double* x, const double* y, const double* z, int32_t rank,
                   int32_t size, fm_handle** out);
void fm_destroy(fm_handle* h);
const char* fm_last_error(const fm_handle* h); /* h may be NULL: error of the last failed fm_create */

/* delaunay_triangulation()  DV:226-356 */
int fm_triangulate(fm_handle* h);
/* Triangulation of the CURRENT generators (after fm_lloyd / fm_set_generators).  warm != 0:
 * keep the previous connectivity and repair it by flip rounds (falls back to a rebuild when
 * a triangle turned over); warm == 0: rebuild from scratch.  No reference equivalent: the
 * reference module is single-shot. */
int fm_retriangulate(fm_handle* h, int32_t warm);
/* sizes for the caller's buffers: T = 2N-4, E = 3N-6 (DV:413, DV:435) */
int fm_counts(const fm_handle* h, int32_t* n_tri, int32_t* n_edge);
/* get_DT: tri and adj are 3 x T column-major (Fortran tri(3,T)) */
int fm_get_triangles(fm_handle* h, int32_t* tri, int32_t* adj);

/* voronoi_diagram(all)  DV:358-585: circumcentres (DV:379-391), cell rings (DV:394-409) and,
 * with all != 0, the Delaunay / Voronoi edge tables (DV:411-460). */
int fm_voronoi(fm_handle* h, int32_t all);
/* get_VVT: xyz 3 x T, lonlat 2 x T (either may be NULL) */
int fm_get_circumcentres(fm_handle* h, double* xyz, double* lonlat);
/* VC%VVT / VC%adjVC rings as CSR: off has N+1 entries (0-based offsets), ring_tri and
 * ring_nbr have 6N-12 entries */
int fm_get_rings(fm_handle* h, int32_t* off, int32_t* ring_tri, int32_t* ring_nbr);
/* get_DE / get_VE (needs fm_voronoi(h, 1)): de 2 x E (DVT1 < DVT2, DV:415-432), ve_vvt 2 x E
 * (the two Voronoi vertices = triangle ids), ve_vc 2 x E (the two cells); edge k of the
 * Voronoi diagram is the dual of Delaunay edge k (DV:435-460). */
int fm_get_edges(fm_handle* h, int32_t* de, int32_t* ve_vvt, int32_t* ve_vc);

/* The arithmetic of the MPAS adapter, on the device: mpas_mesh_adapt (MM:181-345) and
 * calc_weightsOnEdge (MM:442-561).  Needs fm_voronoi(h, 1).  *max_edges receives the largest cell
 * degree (the reference hard-codes maxEdges = 6, MM:19).  strict != 0 keeps the reference's fatal
 * checks (cell / triangle / kite areas each sum to 4 pi within 1e-10, MM:306, :320, :339; the kite
 * fractions of every cell sum to 1, MM:469-475) -- they fire on any obtuse Delaunay triangle.
 * Index conventions are those the reference reaches after its order fix-ups (DV:462-582):
 * ring position j of a cell = verticesOnCell(j), cellsOnCell(j), edgesOnCell(j), edge j running from
 * vertex j to vertex j+1; around a Voronoi vertex counter-clockwise, cellsOnVertex(k) on the left of
 * edgesOnVertex(k), starting at the vertex's lowest edge id.  Edge ids are those of fm_get_edges. */
int fm_mpas(fm_handle* h, int32_t strict, int32_t* max_edges);
/* per edge (E): xyz 3 x E (xEdge,yEdge,zEdge: intersect, SG:316-363), lonlat 2 x E, dvEdge, dcEdge,
 * dv1Edge, dv2Edge, angleEdge (MM:244-266); any pointer may be NULL */
int fm_get_mpas_edges(fm_handle* h, double* xyz, double* lonlat, double* dv, double* dc, double* dv1,
                      double* dv2, double* angle);
/* per cell, in the CSR layout of fm_get_rings (6N-12 entries): edgesOnCell and the kite area of
 * (cell, ring vertex j) -- kiteAreasOnVertex seen from the cell; areaCell has N entries (MM:296-308) */
int fm_get_mpas_cells(fm_handle* h, int32_t* edges_on_cell, double* area_cell, double* kite_on_cell);
/* per Voronoi vertex (T): cellsOnVertex, edgesOnVertex, kiteAreasOnVertex 3 x T; areaTriangle T
 * (MM:228-242, :310-341) */
int fm_get_mpas_vertices(fm_handle* h, int32_t* cells_on_vertex, int32_t* edges_on_vertex,
                         double* area_triangle, double* kite_areas_on_vertex);
/* per edge: nEdgesOnEdge (E), edgesOnEdge and weightsOnEdge, rows of 2 * max_edges (MM:271-294,
 * calc_weightsOnEdge MM:505-556; weights start from zero, which the reference forgets, MM:386/546) */
int fm_get_mpas_edges_on_edge(fm_handle* h, int32_t* n_edges_on_edge, int32_t* edges_on_edge,
                              double* weights_on_edge);

/* Lloyd loop (no reference code).  Runs at most max_iter iterations of
 * triangulate -> circumcentres -> centroids -> generator update, stopping early when the
 * largest chord displacement is <= tol.  warm != 0 repairs the previous triangulation by
 * flip rounds when no triangle turned over, instead of rebuilding it.  Warm iterations are queued
 * in batches (FM_LLOYD_BATCH, default 32) and the host waits once per batch: the convergence
 * test, the "a triangle turned over -> rebuild" decision and the statistics are taken on the
 * device after every iteration, so the result is the same as with one host round trip per
 * iteration -- same iteration count, same *max_disp, bit-identical generators. */
int fm_lloyd(fm_handle* h, const fm_density* rho, int32_t max_iter, double tol, int32_t warm,
             int32_t* iters, double* max_disp);
/* One iteration restricted to this rank's block of cells (fm_create_dist).  The caller
 * then all-gathers the blocks in place (fm_generator_block gives the device pointers) and
 * the next call sees the full set.  *max_disp is the rank-local maximum. */
int fm_lloyd_local(fm_handle* h, const fm_density* rho, int32_t warm, double* max_disp);
/* Device layout for the generator exchange: one array of n_padded 32-byte records
 * { double x, y, z; int64 original_id } in internal (space-filling-curve) order; the rank's
 * block is records [begin, begin + count).  One in-place all-gather of count records per rank
 * completes an iteration. */
int fm_generator_block(fm_handle* h, void** d_records, int64_t* n_padded, int64_t* begin, int64_t* count);
/* Peer exchange over NVLink (sharded handles, one process per GPU of one box): instead of a
 * collective after the Lloyd kernel, the kernel itself stores every updated generator into the
 * generator arrays of all ranks through CUDA IPC peer mappings, between two device-side barriers.
 * fm_peer_handle writes this rank's FM_PEER_HANDLE_BYTES-byte blob (CUDA IPC handle + device
 * UUID); the host exchanges the blobs by whatever means it has (MPI_Allgather,
 * torch.distributed) and passes all `size` of them, in rank order, to fm_peer_attach, which
 * fails if two ranks share a GPU (one GPU per rank is required: the barrier kernels of the ranks
 * wait on each other).  The warm sweep is then sharded as well: each rank tests the triangles of
 * its own cells and sends the edges to flip to all ranks.  Afterwards fm_lloyd runs the whole loop on a sharded handle
 * (every rank calls it with the same arguments) and *max_disp is the maximum over all ranks. */
#define FM_PEER_HANDLE_BYTES 80
int fm_peer_handle(fm_handle* h, void* out80);
int fm_peer_attach(fm_handle* h, const void* handles);
/* generators in input order (device -> host) */
int fm_get_generators(fm_handle* h, double* x, double* y, double* z);
/* replace the generators (host -> device), keeping the internal order and topology.
 * Contract: after fm_set_generators or fm_lloyd (which ends with a centroid update) the stored
 * connectivity is that of the PREVIOUS generators; fm_get_triangles, fm_voronoi and fm_quality
 * repair it first (as fm_retriangulate(h, 1)), so what they return always belongs to the current
 * generators. */
int fm_set_generators(fm_handle* h, const double* x, const double* y, const double* z);

/* One rank's block (fm_create_dist): the host side of rank r holds only the cells it owns.
 * fm_block_ids: the 1-based input-order ids of the block's cells, in block order (count entries,
 * fm_generator_block gives the count; the last rank's block may be shorter: only ids <= N count).
 * fm_get_generators_block / fm_set_generators_block move 3 x count doubles in that order; the
 * upload is stored into EVERY rank's generator array over NVLink (the same record stores the
 * Lloyd kernel uses) and ends with a device-side barrier, so all ranks must call it in step.
 * On an unsharded handle the block is the whole set in internal (space-filling-curve) order. */
int fm_block_ids(fm_handle* h, int32_t* ids);
int fm_get_generators_block(fm_handle* h, double* x, double* y, double* z);
int fm_set_generators_block(fm_handle* h, const double* x, const double* y, const double* z);

/* Mesh-quality statistics of the CURRENT generators: cell-degree histogram
 * (degree d in bin d, d >= 15 in bin 15), Voronoi cell and Delaunay triangle areas by the
 * reference's angle-excess formula (calc_area SG:198-227 over calc_sphere_angle SG:247-285; the sums
 * are its checks MM:306 / MM:320 against 4 pi), smallest / largest interior angle of the Delaunay
 * triangles (radians), and the largest chord between a generator and the centroid of its cell under
 * rho (the displacement the next Lloyd step would make; nothing is moved).  Sums are deterministic. */
typedef struct fm_quality_stats {
  int64_t degree_hist[16];
  double area_cell_min, area_cell_max, area_cell_sum;
  double area_tri_min, area_tri_max, area_tri_sum;
  double angle_min, angle_max;
  double max_disp;
} fm_quality_stats;
int fm_quality(fm_handle* h, const fm_density* rho, fm_quality_stats* out);

/* fm_quality without any allocation.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid it).  Same statistics as fm_quality(h, NULL, out) -- uniform
 * density -- computed by the same kernels; the block partials that fm_quality keeps in a buffer it
 * allocates on first use live in a file-scope device array of the library instead, sized for
 * FM_QUALITY_STATIC_MAX_GENERATORS generators (1.0 MB on every GPU that loads the library, until
 * the process exits).  The call makes no cudaMalloc / cudaFree and no allocating library call.
 * It fails, with fm_last_error saying why, when the handle holds more generators than the limit,
 * when the generators moved since the last triangulation (call fm_retriangulate first: this entry
 * point repairs nothing), or when the handle's device is not the current one.
 * ALL handles and calls share the one array, and it keeps the previous call's contents (every
 * call re-initialises what it reads): calls that overlap on the GPU -- two handles, two host
 * threads -- race on it; serialise them. */
#define FM_QUALITY_STATIC_MAX_GENERATORS 4194304
int fm_quality_static(fm_handle* h, fm_quality_stats* out);

/* Page-locked host buffers for the arrays that cross the boundary (optional: any host
 * pointer is accepted, pinned ones copy at full PCIe rate). */
int fm_host_alloc(size_t bytes, void** out);
int fm_host_free(void* p);

/* Measurement support (bench.py): wall-clock of the device work between begin/end taken
 * with CUDA events on the handle's stream, and per-kernel accounting. */
int fm_timer_begin(fm_handle* h);
int fm_timer_end(fm_handle* h, double* ms);
/* stats[0..7]: insertion rounds, flip sub-rounds, flips, largest flip queue, kernel
 * launches since the last fm_stats_reset, cold rebuilds, warm repairs, cold builds made by stars (fm_star_stats) */
int fm_stats(const fm_handle* h, int64_t* stats);
int fm_stats_reset(fm_handle* h);
/* The cold build "by stars" (every generator's Delaunay neighbours found independently from the points
 * around it, csrc/star_kernels.cuh; replaces the insertion + flip rounds of DV:255-321 / DV:845-882 for
 * well-spaced generators and falls back to them otherwise; FM_COLD_BUILD=flip switches it off).
 * out[0] builds completed that way since the last fm_stats_reset, out[1] attempts that fell back, then of
 * the last attempt: out[2] generators whose near list overflowed (second attempt with long lists), out[3] generators redone with
 * the exact predicates, out[4] stars that
 * did not close, out[5] triangles missing from their owner's star, out[6] triangles, out[7] sum of the
 * degrees, out[8] flips and out[9] turned-over triangles found by the verifying sweep (both must be 0). */
int fm_star_stats(const fm_handle* h, int64_t* out);
/* Per-kernel profile: when enabled every launch is bracketed by CUDA events (slower;
 * never inside a headline timing).  fm_profile_get fills up to cap rows of
 * (name, launches, total_ms, threads) and returns the row count through *rows. */
int fm_profile_enable(fm_handle* h, int32_t on);
int fm_profile_get(fm_handle* h, int32_t cap, char (*names)[32], int64_t* launches, double* ms,
                   int64_t* threads, int32_t* rows);
/* fp64 FMA rate of the device in use, TFLOP/s: independent DFMA chains on every SM, repeated until
 * at least min_ms of device time were measured; the best launch is reported.  The denominator of
 * the fp64-pipe roofline of the centroid / quadrature kernel (no fp64 peak is published for the box). */
int fm_fp64_peak(double min_ms, double* tflops);
/* name and SM count of the device in use */
int fm_device_info(char* name, int32_t cap, int32_t* sms);

#ifdef __cplusplus
}
#endif
#endif
