/* hyteg_b200 -- C-ABI of the B200-native matrix-free P1 geometric-multigrid path.
 *
 * This is the drop-in boundary: plain pointers, sizes and opaque handles, no C++/torch types.  Each
 * entry point names the HyTeG interface it stands in for (paths relative to the reference tree).
 * The C++ header include/hyteg_b200/hyteg.hpp wraps these calls in classes with HyTeG's own names
 * (P1Function, P1ConstantLaplaceOperator, GeometricMultigridSolver, ...); INTEGRATION.md shows the
 * reference-side binding.
 *
 * Conventions: every function returning int returns 0 on success and a non-zero code on failure,
 * with the message available from hb_last_error().  All device work is enqueued on the stream set
 * by hb_set_stream() (default: the legacy default stream); calls that return values to the host
 * synchronise that stream.  DoFType / UpdateType values are HyTeG's (src/hyteg/types/types.hpp:29-44).
 */
#ifndef HYTEG_B200_H
#define HYTEG_B200_H

#ifdef __cplusplus
extern "C" {
#endif

typedef struct hb_storage  hb_storage;  /* PrimitiveStorage (macro mesh + per-level device tables) */
typedef struct hb_function hb_function; /* P1Function<double>                                       */
typedef struct hb_operator hb_operator; /* P1ConstantLaplaceOperator                                */
typedef struct hb_solver   hb_solver;   /* Solver<P1ConstantLaplaceOperator>                        */

enum
{
   HB_None              = 0,
   HB_Inner             = 1,
   HB_DirichletBoundary = 2,
   HB_NeumannBoundary   = 4,
   HB_FreeslipBoundary  = 8,
   HB_Boundary          = 14,
   HB_All               = 15
};
enum
{
   HB_Replace = 0,
   HB_Add     = 1
};
enum
{
   HB_VCYCLE = 0,
   HB_WCYCLE = 1
};

/* ---- runtime ---------------------------------------------------------------------------------- */
const char* hb_last_error( void );
const char* hb_version( void );
int         hb_device_init( int device );         /* cudaSetDevice; fails loudly without a CUDA device */
int         hb_set_stream( void* cuda_stream );    /* cudaStream_t to enqueue all subsequent work on     */
int         hb_synchronize( void );
/* Sum/max reduction across ranks, installed by the host plumbing (NCCL/torch.distributed); stands in
 * for walberla::mpi::allReduceInplace (src/hyteg/p1functionspace/VertexDoFFunction.cpp:1714).
 * op: 0 sum, 2 max.  buf is a HOST buffer of n doubles. */
typedef void ( *hb_allreduce_fn )( double* buf, int n, int op, void* ctx );
int hb_set_allreduce( hb_allreduce_fn fn, void* ctx );
/* ---- ranks: macro-cells are partitioned over n_ranks processes (one per GPU); the ghost layers next to macro primitives
 * shared between ranks are exchanged on demand before every operation that reads them.  Stands in for HyTeG's
 * communicate< From, To >() (src/hyteg/communication/BufferedCommunication.cpp:181-420, VertexDoFPackInfo.hpp).
 * Product path: NCCL send/recv over NVLink on the library's stream.  Rank 0 calls hb_nccl_unique_id, the host plumbing
 * (MPI / torch.distributed) broadcasts the 128 bytes, every rank calls hb_nccl_init; dotGlobal then all-reduces on the
 * device as well.  NCCL is bound at run time (dlopen), the library has no link-time dependency on it. */
int hb_nccl_unique_id( void* id128 );
int hb_nccl_init( const void* id128, int rank, int n_ranks );
int hb_nccl_finalize( void );
/* Peer-memory ghost exchange (one node, NVLink / NVSwitch): on top of the NCCL communicator (still used for reductions),
 * every rank allocates a window (hb_p2p_window returns its 64-byte CUDA IPC handle; segment_bytes bounds one ghost message
 * to one neighbour, 16 bytes per double), the host plumbing all-gathers the handles, hb_p2p_connect maps them.  From then on one kernel per
 * exchange packs into the neighbours' windows, signals, waits and unpacks (no NCCL call on that path).  hb_p2p_active
 * returns 1 when the path is in use and the number of exchanges it has carried. */
int hb_p2p_window( void* handle64, long long segment_bytes, int n_ranks );
int hb_p2p_connect( const void* handles /* n_ranks x 64 bytes, rank order */, int rank, int n_ranks );
int hb_p2p_finalize( void );
int hb_p2p_active( long long* exchanges );
/* Alternative for hosts without NCCL (CPU tests over gloo/MPI): called with the packed send buffers of every neighbouring
 * rank and the receive buffers to fill; buffers live where function data lives (device memory in the product build). */
typedef int ( *hb_exchange_fn )( void* ctx, int n_peers, const int* peers, double* const* send, const long long* n_send,
                                 double* const* recv, const long long* n_recv, void* stream );
int hb_set_exchange( hb_exchange_fn fn, void* ctx );
/* number of ghost exchanges done so far and doubles sent by this rank (all functions, all levels) */
int hb_exchange_stats( long long* calls, long long* doubles_sent );
/* tuning knobs for kernel experiments and A/B tests ("interior", "overlap", "shuffle_depth", "sor_shuffle", "sor_small",
 * "sor_coop", "apply_parts", ...); value INT_MIN restores the built-in default */
int hb_set_tuning( const char* key, int value );

/* ---- PrimitiveStorage  (src/hyteg/primitivestorage/SetupPrimitiveStorage.hpp, PrimitiveStorage.hpp) */
/* cells: 4 vertex ids per macro-tet, in mesh order (MeshInfo).  Macro-cells are distributed over
 * n_ranks in contiguous blocks; this process keeps the cells of 'rank'. */
hb_storage* hb_storage_create( int n_vertices, const double* xyz, int n_cells, const int* cells, int rank, int n_ranks );
/* same with the target rank of every macro-cell given by the caller (SetupPrimitiveStorage::setTargetRank; what the load balancers of
 * src/hyteg/primitivestorage/loadbalancing/SimpleBalancer.hpp:56-78 compute).  cell_rank == NULL: contiguous blocks. */
hb_storage* hb_storage_create_partitioned( int n_vertices, const double* xyz, int n_cells, const int* cells, int rank, int n_ranks, const int* cell_rank );
/* the whole macro mesh of a partitioned storage on one rank (macro-cells in rank-major order, same vertices and mesh boundary flags): the
 * storage of the replicated coarse levels, see hb_solver_create_replicated_coarse_cycle.  No counterpart in the reference, which keeps
 * every level distributed (its coarse-grid agglomeration is left to PETSc, src/hyteg/petsc/PETScLUSolver.hpp). */
hb_storage* hb_storage_create_replica( const hb_storage* st );
void        hb_storage_destroy( hb_storage* st );
/* SetupPrimitiveStorage::setMeshBoundaryFlagsOnBoundary (SetupPrimitiveStorage.cpp:977-988) */
int hb_storage_set_mesh_boundary_flags_on_boundary( hb_storage* st, int flag_on_boundary, int flag_inner, int highest_dim_always_inner );
/* counts[4] = number of macro vertices, edges, faces, cells (global) */
int hb_storage_counts( const hb_storage* st, int* counts );
/* dim 0..3: vertex ids of every primitive of that dimension (1,2,3,4 ints each) */
int hb_storage_primitive_vertices( const hb_storage* st, int dim, int* out );
int hb_storage_get_mesh_boundary_flags( const hb_storage* st, int dim, int* out );
int hb_storage_set_mesh_boundary_flags( hb_storage* st, int dim, const int* in ); /* Primitive::meshBoundaryFlag_ */
int hb_storage_num_local_cells( const hb_storage* st );
int hb_storage_local_cell_global_id( const hb_storage* st, int local_cell );
/* levelinfo::num_microvertices_per_cell (src/hyteg/Levelinfo.hpp:101-105) */
long long hb_cell_array_size( int level );
/* numberOfGlobalDoFs<P1FunctionTag> (src/hyteg/functions/FunctionProperties.hpp) */
long long hb_storage_num_global_dofs( const hb_storage* st, int level );
/* coordinates of every entry of the local arena on 'level': xyz[3 * n_local_cells * cell_array_size] */
int hb_storage_micro_coordinates( const hb_storage* st, int level, double* xyz );

/* neighbouring ranks of this rank and the size of one ghost exchange of a function on 'level' */
int hb_storage_exchange_info( hb_storage* st, int level, int* n_peers, long long* doubles_sent, long long* doubles_received );

/* ---- P1Function  (src/hyteg/p1functionspace/VertexDoFFunction.hpp/.cpp) ------------------------- */
hb_function* hb_p1function_create( hb_storage* st, int min_level, int max_level );
void         hb_p1function_destroy( hb_function* f );
/* BoundaryCondition: meshBoundaryFlag -> DoFType; default create0123BC (0 Inner,1 Dirichlet,2 Neumann,3 Freeslip) */
int hb_p1function_set_boundary_condition( hb_function* f, const int* dof_type_of_flag, int n );
int hb_p1function_get_boundary_condition( const hb_function* f, int* dof_type_of_flag, int n );
int hb_p1function_copy_boundary_condition( hb_function* dst, const hb_function* src ); /* copyBCs */
hb_storage* hb_p1function_storage( const hb_function* f );
int hb_p1function_level_range( const hb_function* f, int* min_level, int* max_level );
/* Whole local arena (n_local_cells * cell_array_size doubles, macro-cell after macro-cell, each in
 * HyTeG's macro-cell layout, src/hyteg/indexing/MacroCellIndexing.hpp:44-52).  Only DoFs whose
 * primitive matches 'flag' are written (interpolate semantics, VertexDoFFunction.cpp interpolate). */
int hb_p1function_upload( hb_function* f, int level, const double* host, int flag );
int hb_p1function_download( const hb_function* f, int level, double* host );
/* stream-ordered copies of the whole local arena without a synchronize (host buffer pinned and kept alive by the caller) */
int hb_p1function_upload_async( hb_function* f, int level, const double* host );
int hb_p1function_download_async( const hb_function* f, int level, double* host );
int hb_p1function_upload_cell( hb_function* f, int level, int local_cell, const double* host );
int hb_p1function_download_cell( const hb_function* f, int level, int local_cell, double* host );
/* replica <- the owned macro-cell arrays of every rank (all-gather through the all-reduce of hb_nccl_init / hb_set_allreduce);
 * owned macro-cell arrays <- their block of the replica (local copy) */
int hb_p1function_gather_to_replica( const hb_function* f, hb_function* replica, int level );
int hb_p1function_scatter_from_replica( hb_function* f, const hb_function* replica, int level );
/* explicit ghost exchange (communicate); operators call it themselves when the mirrored layers are stale */
int hb_p1function_communicate( hb_function* f, int level );
/* raw device pointer of the local arena (for zero-copy interop, e.g. torch tensors) */
double* hb_p1function_device_ptr( hb_function* f, int level );
/* interpolate( constant, level, flag ) */
int hb_p1function_interpolate_constant( hb_function* f, double value, int level, int flag );
/* assign / add: dst (+)= sum_k scalars[k] * srcs[k]   (VertexDoFFunction::assign/add, n <= 4) */
int hb_p1function_assign( hb_function* dst, int n, const double* scalars, hb_function* const* srcs, int level, int flag );
int hb_p1function_add( hb_function* dst, int n, const double* scalars, hb_function* const* srcs, int level, int flag );
int hb_p1function_add_scalar( hb_function* dst, double scalar, int level, int flag );
/* multElementwise: dst = prod_k srcs[k] */
int hb_p1function_mult_elementwise( hb_function* dst, int n, hb_function* const* srcs, int level, int flag );
/* dotGlobal / sumGlobal / getMaxDoFMagnitude (every DoF counted once; all-reduced when a reducer is set) */
int hb_p1function_dot( const hb_function* a, const hb_function* b, int level, int flag, double* out );
int hb_p1function_sum( const hb_function* a, int level, int flag, double* out );
/* Allocation-free variant of hb_p1function_dot.  No counterpart in the reference: added for callers
 * that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).  The per-CTA partial sums live in a
 * file-scope __device__ array of the library (1184 doubles = 9.25 KB, held by every GPU that loads the library until exit) instead of
 * the storage's cudaMalloc'ed buffer; the call allocates nothing, copies nothing to the host and does not synchronise: the result (same
 * kernels, same summation order, bit-identical to hb_p1function_dot) is written to out_device, ONE double in DEVICE memory supplied by the
 * caller, in stream order on the stream of hb_set_stream.  Limits: single-rank storages; at most 2^28 arena entries on the level
 * (n_local_cells * hb_cell_array_size(level)), larger levels are rejected with an error; the level must have been used once before
 * (its tables are built on first use).  All calls share the scratch array, which keeps the last call's partial sums (no slot needs
 * clearing: each call overwrites every slot it reads): calls that overlap on the GPU, i.e. on different streams, race. */
int hb_p1function_dot_no_alloc( const hb_function* a, const hb_function* b, int level, int flag, double* out_device );
int hb_p1function_max_magnitude( const hb_function* a, int level, int flag, double* out );

/* ---- P1ConstantLaplaceOperator  (src/constant_stencil_operator/P1ConstantOperator.hpp,
 *      src/hyteg/p1functionspace/P1Operator.hpp:192-447, src/hyteg/operators/Operator.hpp:76-83) ---- */
hb_operator* hb_p1_constant_laplace_create( hb_storage* st, int min_level, int max_level );
void         hb_operator_destroy( hb_operator* op );
int          hb_operator_level_range( const hb_operator* op, int* min_level, int* max_level );
/* Operator::apply( src, dst, level, flag, updateType ) */
int hb_operator_apply( hb_operator* op, const hb_function* src, hb_function* dst, int level, int flag, int update_type );
/* P1Operator::smooth_jac( dst, rhs, src, omega, level, flag )  (P1Operator.hpp:429-447), one fused kernel */
int hb_operator_smooth_jac( hb_operator* op, hb_function* dst, const hb_function* rhs, const hb_function* src, double omega, int level, int flag );
/* P1Operator::smooth_sor / smooth_sor_backwards / smooth_gs (P1Operator.hpp:322-427): multicolour ordering */
int hb_operator_smooth_sor( hb_operator* op, hb_function* dst, const hb_function* rhs, double omega, int level, int flag, int backwards );
/* same with the interior ordering named by the caller.  Both 0 and 1 currently select the four-colour ordering (x + 2y + 3z) mod 4 on
 * every level; a forward and a backward sweep of it are exact transposes of each other (SymmetricGaussSeidelSmoother.hpp pairs them). */
int hb_operator_smooth_sor_ordered( hb_operator* op, hb_function* dst, const hb_function* rhs, double omega, int level, int flag, int backwards, int ordering );
/* fused r = b - A x  (GeometricMultigridSolver.hpp:255-259 does apply + assign) */
int hb_operator_residual( hb_operator* op, const hb_function* x, const hb_function* b, hb_function* r, int level, int flag );
/* fused Chebyshev building blocks (ChebyshevSmoother.hpp:179-222):
 *   first: dst = D^-1 (b - A x);   step: dst = D^-1 A src, x += coeff * src */
int hb_operator_cheb_first( hb_operator* op, const hb_function* x, const hb_function* b, hb_function* dst, int level, int flag );
int hb_operator_cheb_step( hb_operator* op, const hb_function* src, hb_function* dst, hb_function* x, double coeff, int level, int flag );
/* computeDiagonalOperatorValues / computeInverseDiagonalOperatorValues (P1Operator.hpp:452-465) into f */
int hb_operator_diagonal( hb_operator* op, hb_function* f, int level, int inverse );
/* the 15 macro-cell stencil weights of a local macro-cell for boundary mask 'mask' (0 = interior), order:
 * C W E N S NW SE TC TW TS TSE BC BN BE BNW  (P1ConstantOperator cell stencil, P1Elements.hpp) */
int hb_operator_cell_stencil( const hb_operator* op, int level, int local_cell, int mask, double* w15 );

/* ---- P1toP1LinearRestriction / P1toP1LinearProlongation (src/hyteg/gridtransferoperators/) ------- */
int hb_p1_restrict( hb_function* f, int source_level, int flag );            /* restrict( f, sourceLevel, flag ) */
int hb_p1_prolongate( hb_function* f, int source_level, int flag );          /* prolongate( f, sourceLevel, flag ) */
int hb_p1_prolongate_and_add( hb_function* f, int source_level, int flag );  /* prolongateAndAdd */

/* ---- Solvers (src/hyteg/solvers/) ------------------------------------------------------------------ */
hb_solver* hb_solver_create_cg( hb_storage* st, int min_level, int max_level, int max_iter, double rel_tol, double abs_tol ); /* CGSolver.hpp:44-91 */
hb_solver* hb_solver_create_weighted_jacobi( hb_storage* st, int min_level, int max_level, double omega ); /* WeightedJacobiSmoother.hpp */
hb_solver* hb_solver_create_sor( double omega );                                                            /* SORSmoother.hpp */
hb_solver* hb_solver_create_gauss_seidel( void );                                                           /* GaussSeidelSmoother.hpp */
hb_solver* hb_solver_create_symmetric_gauss_seidel( void );                                                 /* SymmetricGaussSeidelSmoother.hpp */
hb_solver* hb_solver_create_chebyshev( hb_storage* st, int min_level, int max_level );                      /* ChebyshevSmoother.hpp */
/* ChebyshevSmoother::setupCoefficients( order, spectralRadius ) / ( order, lower, upper ) (:240-276) */
int hb_chebyshev_setup_coefficients( hb_solver* cheb, int order, double lower_bound, double upper_bound );
int hb_chebyshev_coefficients( const hb_solver* cheb, double* coeffs, int* n );
/* chebyshev::estimateRadius (ChebyshevSmoother.hpp:637-647 -> numerictools/SpectrumEstimation.hpp:56-85):
 * power iteration on D^-1 A starting from x (overwritten) */
int hb_estimate_spectral_radius( hb_operator* op, hb_function* x, hb_function* tmp, int iterations, int level, double* radius );
/* GeometricMultigridSolver (GeometricMultigridSolver.hpp:61-96) with P1 linear transfer operators */
hb_solver* hb_solver_create_gmg( hb_storage* st, hb_solver* smoother, hb_solver* coarse_solver, int min_level, int max_level,
                                 int pre_smooth, int post_smooth, int smooth_increment, int cycle_type );
/* Coarse-grid solver for a GeometricMultigridSolver on levels coarse_level .. L of a PARTITIONED storage: gathers x and b of level
 * coarse_level onto a replica of the whole mesh and runs there, on every rank redundantly, the rest of the cycle -- V(pre,post) on levels
 * 0 .. coarse_level with smoother 0 = Gauss-Seidel / 1 = weighted Jacobi( omega ) and CG( max_iter, rel_tol, abs_tol ) on level 0 -- i.e.
 * the same operations GeometricMultigridSolver::solveRecursively (GeometricMultigridSolver.hpp:140-195) performs on those levels, without
 * a single ghost exchange.  The macro-cell numbering of the replica differs from the partitioned one, so results agree with the fully
 * distributed cycle up to rounding, not bitwise.  cg_max_iter == 0: level 0 is solved directly (hb_solver_create_direct_level0). */
hb_solver* hb_solver_create_replicated_coarse_cycle( hb_storage* st, int coarse_level, int smoother, double omega, int pre, int post, int cg_max_iter,
                                                     double cg_rel_tol, double cg_abs_tol );
/* Direct solver for level 0 of an unpartitioned storage (the role PETScLUSolver plays as a coarse-grid solver in the reference,
 * src/hyteg/petsc/PETScLUSolver.hpp): dense Cholesky of the free macro-vertex DoFs, assembled once per boundary-condition table by applying
 * the operator to unit vectors; a solve is one download, two triangular sweeps on the host and one upload. */
hb_solver* hb_solver_create_direct_level0( hb_storage* st );
int        hb_solver_solve( hb_solver* s, hb_operator* op, hb_function* x, hb_function* b, int level ); /* Solver::solve */
int        hb_solver_iterations( const hb_solver* s );
void       hb_solver_destroy( hb_solver* s );

/* ==== P2 function space, geometry blending and the fused elementwise operators (BASELINE configs 3 and 4) ============================
 * One rank only.  Data layout: the vertex DoFs are a P1Function (hb_p2function_vertex_part: all hb_p1function_* calls apply); the edge
 * DoFs of a macro-cell are HyTeG's seven orientation arrays X Y Z XY XZ YZ (width 2^L) and XYZ (width 2^L - 1) one behind the other
 * (src/hyteg/edgedofspace/EdgeDoFIndexing.hpp:922-985), macro-cell after macro-cell, covering the closed macro-tet. */
typedef struct hb_p2function hb_p2function; /* P2Function<double>  (src/hyteg/p2functionspace/P2Function.hpp)                         */
typedef struct hb_p2operator hb_p2operator; /* P2ElementwiseOperator<Form> (src/hyteg/elementwiseoperators/P2ElementwiseOperator.hpp)  */

/* IcosahedralShellMap::setMap( setupStorage ) (src/hyteg/geometry/IcosahedralShellMap.hpp:491-510): every macro-cell gets the shell map
 * of its prism (classifyVertices, :608-732).  Call before creating P2 operators. */
int hb_storage_set_icosahedral_shell_map( hb_storage* st );
/* rayVertex, refVertex, thrVertex, forVertex (3 doubles each), radRefVertex, radRayVertex of a macro-cell's map (:512-518) */
int hb_storage_shell_map_parameters( hb_storage* st, int local_cell, double* out14 );
/* GeometryMap::evalF / evalDF (:109-147, :169-239) on the device at n points x[n][3] of the macro-cell: out[n][12] = F (3), DF (9, row-major) */
int hb_storage_map_eval( hb_storage* st, int level, int local_cell, const double* x, int n, double* out );

hb_p2function* hb_p2function_create( hb_storage* st, int min_level, int max_level );
void           hb_p2function_destroy( hb_p2function* f );
hb_function*   hb_p2function_vertex_part( hb_p2function* f ); /* P2Function::getVertexDoFFunction(); owned by f */
long long      hb_p2_edge_cell_size( int level );             /* entries of one macro-cell's edge block */
int            hb_p2function_upload_edges( hb_p2function* f, int level, const double* host );   /* P2Function::getEdgeDoFFunction() data */
int            hb_p2function_download_edges( const hb_p2function* f, int level, double* host );
double*        hb_p2function_edge_device_ptr( hb_p2function* f, int level );
/* computational coordinates of the two end points of every entry of a macro-cell's edge block: xyz[entries][2][3] */
int hb_p2_edge_endpoints( hb_storage* st, int level, int local_cell, double* xyz );
/* P2Function::interpolate( constant ) / assign / add / dotGlobal (src/hyteg/p2functionspace/P2Function.cpp) */
int hb_p2function_interpolate_constant( hb_p2function* f, double value, int level, int flag );
int hb_p2function_assign( hb_p2function* dst, int n, const double* scalars, hb_p2function* const* srcs, int level, int flag );
int hb_p2function_add( hb_p2function* dst, int n, const double* scalars, hb_p2function* const* srcs, int level, int flag );
int hb_p2function_dot( const hb_p2function* a, const hb_p2function* b, int level, int flag, double* out );
int hb_p2function_mult_elementwise( hb_p2function* dst, int n, hb_p2function* const* srcs, int level, int flag );
int hb_p2function_copy_boundary_condition( hb_p2function* dst, const hb_p2function* src );

/* P2ElementwiseOperator< forms::p2_div_k_grad_blending_q4 > (P2ElementwiseBlendingDivKGradOperator, P2ElementwiseOperator.hpp:475) /
 * operatorgeneration::P2ElementwiseDivKGrad...: k as a P2Function (the hyteg_operators way; null: k = 1), or sampled at the quadrature
 * points from a host callback (the in-tree form's std::function): hb_p2operator_quadrature_points tells where, in the order
 * [macro-cell][micro-cell type][micro-cell index][quadrature point]; entries of micro-cells that do not exist are (0,0,0) / ignored. */
hb_p2operator* hb_p2_divkgrad_create( hb_storage* st, int min_level, int max_level, const hb_p2function* k );
/* P2ElementwiseOperator< forms::p2_epsilonvar_i_j_blending_q2 >: block (i, j) of P2ElementwiseBlendingEpsilonOperator
 * (src/hyteg/elementwiseoperators/P2ElementwiseEpsilonOperator.hpp:112-170), viscosity mu like k above */
hb_p2operator* hb_p2_epsilon_create( hb_storage* st, int min_level, int max_level, int i, int j, const hb_p2function* mu );
void           hb_p2operator_destroy( hb_p2operator* op );
long long      hb_p2operator_num_quadrature_points( hb_p2operator* op, int level );
int            hb_p2operator_quadrature_points( hb_p2operator* op, int level, double* xyz ); /* physical (blended) positions */
int            hb_p2operator_set_sampled_coefficient( hb_p2operator* op, int level, const double* values );
/* Operator::apply( src, dst, level, flag, updateType ) -- P2ElementwiseOperator::gemv (P2ElementwiseOperator.cpp:109-330) */
int hb_p2operator_apply( hb_p2operator* op, const hb_p2function* src, hb_p2function* dst, int level, int flag, int update_type );

/* computeDiagonalOperatorValues / computeInverseDiagonalOperatorValues (P2ElementwiseOperator.cpp:332-420): f has one entry, three
 * for the vector epsilon operator */
int hb_p2operator_diagonal( hb_p2operator* op, hb_p2function* const* f, int level, int inverse );
/* P2ToP1ElementwiseBlendingDiv{x,y,z}Operator (P2ToP1ElementwiseOperator.hpp:264-266, forms::p2_to_p1_div_c_blending_q2) and
 * P1ToP2ElementwiseBlendingDivT{x,y,z}Operator (P1ToP2ElementwiseOperator.hpp:275-277, forms::p1_to_p2_divt_c_blending_q2): the B and
 * B^T blocks of P2P1ElementwiseBlendingStokesOperator (P2P1ElementwiseBlendingStokesOperator.hpp:91-100); component 0..2 */
hb_p2operator* hb_p2_div_create( hb_storage* st, int min_level, int max_level, int component );
hb_p2operator* hb_p2_divt_create( hb_storage* st, int min_level, int max_level, int component );
int hb_p2operator_apply_p2_to_p1( hb_p2operator* op, const hb_p2function* src, hb_function* dst, int level, int flag, int update_type );
int hb_p2operator_apply_p1_to_p2( hb_p2operator* op, const hb_function* src, hb_p2function* dst, int level, int flag, int update_type );
/* P2ElementwiseBlendingEpsilonOperator (P2ElementwiseEpsilonOperator.hpp:112-170) as ONE fused kernel: all nine blocks on a P2 vector
 * function (three P2Functions), the geometry of a quadrature point evaluated once */
hb_p2operator* hb_p2_epsilon_vector_create( hb_storage* st, int min_level, int max_level, const hb_p2function* mu );
int hb_p2operator_apply_vector( hb_p2operator* op, hb_p2function* const* src3, hb_p2function* const* dst3, int level, int flag, int update_type );

/* The whole P2-P1 Taylor-Hood Stokes operator with the epsilon velocity block in ONE fused kernel: dst.uvw = A src.uvw + B^T src.p,
 * dst.p = B src.uvw (P2P1ElementwiseBlendingStokesOperator::apply, P2P1ElementwiseBlendingStokesOperator.hpp:91-100, with
 * P2ElementwiseBlendingEpsilonOperator as A: src/hyteg_operators_composites/stokes/P2P1StokesEpsilonOperator.hpp:50-73); Replace only */
hb_p2operator* hb_p2_stokes_create( hb_storage* st, int min_level, int max_level, const hb_p2function* mu );
int hb_p2operator_apply_stokes( hb_p2operator* op, hb_p2function* const* src_uvw, const hb_function* src_p, hb_p2function* const* dst_uvw, hb_function* dst_p,
                                int level, int flag );

/* Chebyshev-smoothed geometric multigrid on a P2 elementwise operator (BASELINE config 4; the velocity-block recipe of
 * src/terraneo/solvers/StokesMCFGMRESSolver.hpp:71-143): GeometricMultigridSolver + ChebyshevSmoother with the inverse diagonal
 * (radius from power iterations started at eigenvector_guess on max_level, ChebyshevSmoother.hpp:655-665) + quadratic transfer, CG on
 * min_level.  what = 0: one V(pre,post) cycle, 1: one call of the smoother. */
typedef struct hb_p2solver hb_p2solver;
hb_p2solver* hb_p2_gmg_chebyshev_create( hb_storage* st, hb_p2operator* op, int min_level, int max_level, int order, hb_p2function* eigenvector_guess,
                                         int power_iterations, int pre, int post, int coarse_max_iter, double coarse_tol );
double       hb_p2solver_spectral_radius( const hb_p2solver* s );
int          hb_p2solver_solve( hb_p2solver* s, hb_p2function* x, hb_p2function* b, int level, int what );
void         hb_p2solver_destroy( hb_p2solver* s );

/* P2toP2QuadraticRestriction::restrict / P2toP2QuadraticProlongation::prolongate / prolongateAndAdd
 * (src/hyteg/gridtransferoperators/P2toP2QuadraticRestriction.cpp:131-251, P2toP2QuadraticProlongation.cpp) */
int hb_p2_restrict( hb_p2function* f, int source_level, int flag );
int hb_p2_prolongate( hb_p2function* f, int source_level, int flag );
int hb_p2_prolongate_and_add( hb_p2function* f, int source_level, int flag );

#ifdef __cplusplus
}
#endif
#endif
