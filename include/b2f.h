/*
 * b2f.h — C ABI of libb200fem.so: B200-native (sm_100a) finite-element bilinear-form assembly.
 *
 * Drop-in boundary for ArcaneFEM's femutils assembly hot path (reference: toutane/arcanefem; all
 * file:line citations are relative to the reference root).  In the reference this path is not an FFI
 * but a set of header-only templates compiled with each module:
 *   femutils/BSRFormat.h              BSRMatrix<BS> / BSRFormat<NB_DOF>
 *   femutils/ArcaneFemFunctionsGpu.h  element helpers, RHS and Dirichlet flagging kernels
 *   femutils/HypreDoFLinearSystem.cc  _applyRowElimination / _applyForcedValuesToLhs (:319-382)
 *   femutils/DoFLinearSystem.h        CSRFormatView hand-off (:42-76)
 * A C ABI cannot carry a device lambda, so the reference's `Function compute_element_matrix` template
 * parameter (BSRFormat.h:786-787,1105) becomes an operator enum + parameter block (b2f_operator); the
 * header-only C++ shim in include/arcanefem_b200/ restores the reference's names on top of this file and
 * include/b2f/assemble.cuh keeps a template entry point for user-written element functors.
 *
 * Conventions: every function returns B2F_OK (0) or a negative b2f_status; no exception crosses the
 * boundary; b2f_last_error_string() describes the last failure on the calling thread.  All data pointers
 * are device (or managed) pointers unless the name says host.  `stream` is a cudaStream_t passed as
 * void*.  Calls on one handle are not thread-safe and are ordered on the stream given.  The library
 * has no CPU fallback: without a CUDA device every compute entry point fails with B2F_ERR_CUDA.
 */
#ifndef B2F_H
#define B2F_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B2F_VERSION_MAJOR 0
#define B2F_VERSION_MINOR 3

typedef struct b2f_context_s* b2f_handle;
typedef void* b2f_stream; /* cudaStream_t */

typedef enum b2f_status {
  B2F_OK = 0,
  B2F_ERR_INVALID_ARGUMENT = -1, /* reference: ARCANE_THROW(ArgumentException), BSRFormat.h:121-125 */
  B2F_ERR_NOT_IMPLEMENTED = -2, /* reference: NotImplementedException, BSRFormat.h:393-394 */
  B2F_ERR_CUDA = -3,
  B2F_ERR_STATE = -4, /* call order violated (e.g. assemble before compute_sparsity) */
  B2F_ERR_NOT_FOUND = -5, /* reference: std::runtime_error("Value not found"), BSRFormat.h:170 */
  B2F_ERR_OVERFLOW = -6, /* a per-rank count does not fit the reference's Int32 indices */
  B2F_ERR_NCCL = -7
} b2f_status;

/* Value layout, femutils/BSRFormat.h:145-171 (findValueIndex) and :405. */
typedef enum b2f_layout {
  B2F_LAYOUT_PER_BLOCK = 0, /* order_values_per_block = true  (Aleph path): p*BS*BS + i*BS + j           */
  B2F_LAYOUT_PER_ROW = 1 /* order_values_per_block = false (Hypre/CSR): row_index[r]*BS*BS + i*BS*k + x*BS + j */
} b2f_layout;

/* Scatter-add variant, femutils/BSRFormat.h:1105-1111. */
typedef enum b2f_assembly_variant {
  B2F_ASM_DEFAULT = 0, /* the library's choice (node gather) */
  B2F_ASM_CELL_ATOMIC = 1, /* cell-wise, one fp64 atomic per entry (BSRFormat.h:786-898) */
  B2F_ASM_CELL_ATOMIC_WARPAGG = 2, /* cell-wise, same-address adds pre-reduced inside the warp */
  B2F_ASM_NODE_GATHER = 3, /* node-wise, atomic-free, deterministic (BSRFormat.h:937-1068); fan-walk plan */
  B2F_ASM_NODE_GATHER_SLOTS = 4 /* first-generation node gather (one column-slot word per node/cell pair); kept for A/B measurements */
} b2f_assembly_variant;

typedef enum b2f_operator_kind {
  B2F_OP_LAPLACE = 0, /* modules/poisson/FemModule.h:139-147 (tri), :177-186 (tet); BS = 1 */
  B2F_OP_ELASTICITY = 1, /* modules/elasticity/FemModule.h:112-152 (tri, BS = 2); tet BS = 3 by extension */
  B2F_OP_HEAT = 2 /* modules/heat/FemModule.cc:678-770 (tri, BS = 1): lambda*stiffness + mass/dt */
} b2f_operator_kind;

/* Dirichlet elimination modes stored in elimination_info (HypreDoFLinearSystem.cc:237-241,
 * AlephDoFLinearSystem.cc ELIMINATE_ROW_COLUMN). */
enum { B2F_ELIMINATE_NONE = 0, B2F_ELIMINATE_ROW = 1, B2F_ELIMINATE_ROW_COLUMN = 2 };

/*
 * Mesh as the reference's kernels see it: cellNode() connectivity (fixed arity, Int32 local ids),
 * VariableNodeReal3 coordinates (array of 3 x fp64, z present even in 2D), isOwn(node) flags
 * (femutils/BSRFormat.h:789-792).  The library never frees these pointers and reads them at every
 * call that needs them, so they must stay valid while the handle is used.
 */
typedef struct b2f_mesh_view {
  int32_t dim; /* 2 or 3 */
  int32_t nodes_per_cell; /* 3 (Triangle3) or 4 (Tetraedron4) */
  int64_t nb_cell;
  int64_t nb_node;
  const int32_t* cell_node; /* [nb_cell * nodes_per_cell] */
  const double* coords; /* [nb_node * 3], xyz interleaved */
  const uint8_t* node_is_own; /* [nb_node] or NULL = every node is own */
  const uint8_t* cell_is_own; /* [nb_cell] or NULL = every cell is assembled; 0 = ghost cell: pattern only (multi-GPU exchange mode) */
} b2f_mesh_view;

/* Walk of the node-gather kernel over the cells around a node (tets).  SELECT (what AUTO picks): one step body, the
 * roles of the window vertices are chosen with selects.  LOCKSTEP (kept for A/B measurements): the step word names one
 * of six select-free step bodies, reached through an indirect branch; rows of a warp in different bodies are serialised
 * (stats.fan_lockstep_bodies_milli).  Measured slower than SELECT even on structured meshes, see DESIGN.md. */
typedef enum b2f_fan_mode { B2F_FAN_AUTO = 0, B2F_FAN_SELECT = 1, B2F_FAN_LOCKSTEP = 2 } b2f_fan_mode;

/* How b2f_compute_sparsity forms the pattern.  SORT: the reference's algorithm (femutils/BSRFormat.h:602-744): cell -> edge
 * keys, global radix sort, run ends, degree scan, column fill.  LISTS: node -> cell lists first, then every row gathers
 * the vertices of its cells into a sorted set - no edge keys and no global sort, same pattern bit for bit.  AUTO = LISTS
 * (in the cell_is_own exchange mode the lists of the pattern hold the ghost cells too and the lists of the assembled cells
 * are built in a second pass), except for rows with more than 64 neighbours, which take SORT. */
typedef enum b2f_sparsity_build { B2F_SPARSITY_AUTO = 0, B2F_SPARSITY_SORT = 1, B2F_SPARSITY_LISTS = 2 } b2f_sparsity_build;

/* Order of the blocks inside a row.  DIAG_FIRST (default): diagonal, then the neighbours by ascending id - what the sorted-edge
 * build of the reference gives when executed in key order (femutils/BSRFormat.h:602-744).  DIAG_LAST: the neighbours first, the
 * diagonal last - the layout of the reference's connectivity ("atomic-free") sparsity, computeSparsityAtomicFree :479-578, which
 * lists a node's neighbours in the order of Arcane's node-node / node-face connectivity and then appends the node itself.  That
 * connectivity order is a property of Arcane's edge numbering and is not reproducible outside Arcane: the
 * neighbours are listed by ascending id here; the per-row neighbour SET is the same as DIAG_FIRST's. */
typedef enum b2f_pattern_layout { B2F_PATTERN_DIAG_FIRST = 0, B2F_PATTERN_DIAG_LAST = 1 } b2f_pattern_layout;

/* Node locality inside the library.  The node-gather kernel walks 128 consecutive rows per CTA and gathers the coordinates of
 * their neighbours: it is fast when consecutive node ids are close in space (any mesh generator's numbering, an RCM or a
 * space-filling order) and several times slower when they are not (a random renumbering: 6.8 ms instead of 0.95 ms on the
 * 50 M-tet cube).  The reference's one-thread-per-cell kernels have no such precondition (femutils/BSRFormat.h:807-836), so the
 * library can supply the locality itself: b2f_compute_sparsity sorts the nodes along a space-filling order of their
 * coordinates, renumbers the connectivity INTERNALLY and builds the kernel's plan (tiles, column ids, walk words) in that
 * numbering; every assembly permutes the coordinates into it (one streaming pass) and the kernel writes each row's values
 * where the CALLER's row_index and column order put them.  Pattern, values, right-hand side and every view stay in the
 * caller's numbering (values agree with the unreordered kernel to rounding: the rows are walked in another order).
 * Applies to the node-gather variant on a single rank (ignored with cell_is_own and for the other variants). */
typedef enum b2f_reorder {
  B2F_REORDER_NONE = 0,
  B2F_REORDER_MORTON = 1, /* Morton (Z) curve of the coordinates: compact 3D tiles */
  B2F_REORDER_PENCIL = 2 /* bins of one mean node spacing in y and z, nodes sorted by x inside a bin: a tile is a run along x, like the
                          * lexicographic numbering of a structured mesh, so the lanes of a warp gather neighbouring coordinates */
} b2f_reorder;

/* Mirrors BSRFormat<NB_DOF>::initialize(mesh, does_linear_system_use_csr, use_atomic_free), BSRFormat.h:389. */
typedef struct b2f_options {
  int32_t block_size; /* NB_DOF: 1, 2 or 3 */
  int32_t use_csr; /* does_linear_system_use_csr: 1 -> B2F_LAYOUT_PER_ROW, 0 -> B2F_LAYOUT_PER_BLOCK */
  int32_t assembly_variant; /* b2f_assembly_variant; the reference's use_atomic_free=true maps to NODE_GATHER */
  int32_t fan_mode; /* b2f_fan_mode: walk of the node-gather kernel on tets (0 = chosen from the plan) */
  int32_t sparsity_build; /* b2f_sparsity_build */
  int32_t pattern_layout; /* b2f_pattern_layout */
  int32_t reorder; /* b2f_reorder */
  int32_t reserved;
} b2f_options;

typedef struct b2f_operator {
  int32_t kind; /* b2f_operator_kind */
  int32_t accumulate; /* 0: values are overwritten (zero-fill + add); 1: added to current values */
  double lambda; /* elasticity: Lame lambda; heat: conductivity when cell_lambda == NULL */
  double mu2; /* elasticity: 2*mu (modules/elasticity/FemModule.cc:133) */
  double inv_dt; /* heat: 1/dt */
  const double* cell_lambda; /* heat: per-cell conductivity [nb_cell] or NULL */
} b2f_operator;

/* BSRMatrix accessors, femutils/BSRFormat.h:312-320.  row_index has nb_row+1 entries (the reference
 * exposes the first nb_row, :138; entry nb_row equals nb_col). */
typedef struct b2f_bsr_view {
  int32_t block_size;
  int32_t layout; /* b2f_layout */
  int64_t nb_row; /* block rows = nb_node */
  int64_t nb_col; /* stored blocks = 2*nb_edge + nb_node (BSRFormat.h:401) */
  int64_t nb_nz; /* scalars = BS*BS*nb_col (BSRFormat.h:402) */
  int64_t nb_edge;
  const int32_t* row_index;
  const int32_t* columns;
  const int32_t* nb_nz_per_row;
  double* values;
} b2f_bsr_view;

/* CSRFormatView, femutils/DoFLinearSystem.h:42-76; rows has nb_row+1 entries (reference: nb_row). */
typedef struct b2f_csr_view {
  int64_t nb_row; /* BS * nb_node */
  int64_t nb_value;
  const int32_t* rows;
  const int32_t* rows_nb_column;
  const int32_t* columns;
  double* values;
} b2f_csr_view;

/* Per-DoF state of the linear system the Dirichlet/RHS kernels write
 * (femutils/HypreDoFLinearSystem.cc:224-242); dof = node*BS + i (femutils/FemDoFsOnNodes.cc:79-109). */
typedef struct b2f_system_view {
  int64_t nb_dof;
  double* rhs;
  uint8_t* forced_info;
  double* forced_value;
  uint8_t* elimination_info;
  double* elimination_value;
} b2f_system_view;

/* Neumann group, femutils/ArcaneFemFunctionsGpu.h:621-677 (2D, faces = edges) and :711-768 (3D). */
typedef struct b2f_neumann {
  int64_t nb_face;
  const int32_t* face_node; /* [nb_face * (dim)] node ids of each boundary face */
  const uint8_t* face_outside; /* isSubDomainBoundaryOutside per face, or NULL = all outside */
  int32_t has_value; /* 1: scalar `value`; 0: flux vector (value_x, value_y, value_z) . normal */
  int32_t reserved;
  double value;
  double value_x, value_y, value_z;
} b2f_neumann;

/* Plan of the node-gather kernels (library-internal layout, see arcanefem_b200/csrc/sparsity.cu). */
typedef struct b2f_plan_view {
  int64_t nb_node;
  int64_t nb_pair; /* node/cell incidences */
  int64_t nb_fan_step;
  const int32_t* nc_ptr; /* [nb_node + 1] */
  const uint32_t* nc_list; /* [nb_pair] cell * 4 + local index of the node in the cell, ascending per node */
  const int32_t* fan_ptr; /* [nb_node + 1] or NULL when the variant has no fan plan */
  const uint32_t* fan_steps; /* [nb_fan_step] */
  const uint32_t* fan_cell; /* [nb_fan_step] */
} b2f_plan_view;

typedef struct b2f_stats {
  int64_t nb_edge;
  int64_t nb_col;
  int64_t nb_nz;
  int64_t columns_not_found; /* (row,col) pairs the last assemble could not place (reference: silently skipped) */
  int64_t workspace_bytes; /* temporary device bytes held (sort buffers); 0 after b2f_release_workspace */
  int64_t resident_bytes; /* device bytes owned by the handle, workspace excluded */
  int64_t gather_fallback_tiles; /* node-gather tiles that overflowed shared memory and ran on the global path */
  int64_t fan_steps; /* words of the fan-walk plan (incident node/cell pairs + chain starts) */
  int64_t sparsity_by_lists; /* 1: the pattern came from the row-wise build (b2f_sparsity_build) */
  int64_t fan_lockstep; /* 1: the plan is in the lockstep form (b2f_fan_mode) */
  int64_t fan_lockstep_bodies_milli; /* 1000 x step bodies per warp step the lockstep walk executes on this plan (1000 = rows in step) */
  int64_t reordered; /* 1: the plan is stored in the library's own node order (b2f_reorder) */
} b2f_stats;

int b2f_version(void);
const char* b2f_last_error_string(void);
const char* b2f_status_string(int status);

/* BSRFormat ctor + initialize (BSRFormat.h:358-363,389-409).  Sizes are not needed up front: the edge
 * count the reference computes on the host (:370-385) is a by-product of b2f_compute_sparsity. */
int b2f_create(const b2f_mesh_view* mesh, const b2f_options* options, b2f_handle* out);
int b2f_destroy(b2f_handle h);
/* Re-point the handle at new coordinate / ownership arrays (same connectivity). */
int b2f_set_coords(b2f_handle h, const double* coords);

/* computeSparsity (BSRFormat.h:775-781): K1-K6, canonical pattern = diagonal first, then ascending
 * neighbour ids (what :602-744 produces when executed in key order).  Also builds the node->cell lists
 * used by the node-gather variant.  Synchronises `stream` once (the block count must reach the host). */
int b2f_compute_sparsity(b2f_handle h, b2f_stream stream);
int b2f_release_workspace(b2f_handle h);

/* assembleBilinear (BSRFormat.h:1105-1111). */
int b2f_assemble(b2f_handle h, const b2f_operator* op, b2f_stream stream);
/* assembleBilinear and BoundaryConditions{2D,3D}::applyConstantSourceToRhs (ArcaneFemFunctionsGpu.h:402-429,605,695) in
 * one pass: rhs[dof(n,0)] += qdot * |cell| / Nv over the cells around every own node.  With the node-gather variant the
 * cell measures are summed along the same walk that forms the matrix row (the determinant is already there): no atomics,
 * no second sweep over the cells, bitwise reproducible.  Other variants (and the cell_is_own exchange mode, where a
 * rank only walks its own cells) run b2f_assemble followed by b2f_rhs_constant_source. */
int b2f_assemble_with_source(b2f_handle h, const b2f_operator* op, double qdot, b2f_stream stream);
/* User-written element operators: the C ABI cannot carry the reference's `template <class Function> assembleBilinear(Function)`
 * (BSRFormat.h:1105-1111), so a caller that has its own element callback compiles include/b2f/user_operator.cuh with nvcc
 * and calls b2f::assemble_with(handle, op, ...) there; that template fetches the kernel argument block of the handle through
 * this entry (args_out = b2f::AssembleArgs*, args_size = sizeof of the caller's copy; a mismatch of it, of the arity or of the
 * block size is refused) and launches the
 * same kernels the built-in operators use.  Clears the handle's error counters on `stream`. */
int b2f_user_assemble_begin(b2f_handle h, int32_t nodes_per_cell, int32_t block_size, void* args_out, size_t args_size, int32_t* variant_out,
                            int32_t* per_block_out, int64_t* nb_values_out, b2f_stream stream);
/* resetMatrixValues (BSRFormat.h:1120-1123). */
int b2f_reset_values(b2f_handle h, b2f_stream stream);

int b2f_get_bsr(b2f_handle h, b2f_bsr_view* out);
/* BSRMatrix::toCsr + CSRFormatView (BSRFormat.h:194-256,414-430): zero-copy alias of the BSR arrays when
 * BS = 1; device expansion of rows/columns when BS > 1 (values alias; needs B2F_LAYOUT_PER_ROW). */
int b2f_to_csr(b2f_handle h, b2f_csr_view* out, b2f_stream stream);
/* Hand-off helper for a distributed solver (femutils/HypreDoFLinearSystem.cc:474-519, done there in host loops over
 * every row and every entry): rows_global_out[i] = dof_number[i] for the nb_dof local rows and
 * cols_global_out[q] = dof_number[csr.columns[q]] for the nb_value entries of the CSR view, on the device.
 * dof_number is the device array of the global matrix numbering of the local dofs (own: first_own_row + own index,
 * ghosts: the owner's number, :264-303).  Either output may be NULL. */
int b2f_csr_global_numbering(b2f_handle h, const int32_t* dof_number, int32_t* rows_global_out, int32_t* cols_global_out, b2f_stream stream);
/* The argument arrays of HYPRE_IJMatrixSetValues(ij_A, nb_row, ncols, rows, cols, values) and of
 * HYPRE_IJVectorSetValues(ij_b, nb_row, rows, rhs) (femutils/HypreDoFLinearSystem.cc:587-640), device resident:
 * only the rows of own nodes, in local order, with global row / column ids.  Borrowed; valid until the next
 * b2f_to_hypre_ij / b2f_compute_sparsity / b2f_destroy. */
typedef struct b2f_ij_view {
  int64_t nb_row; /* own scalar rows */
  int64_t nb_value; /* entries of those rows */
  const int32_t* ncols; /* [nb_row] row lengths */
  const int32_t* rows; /* [nb_row] global row ids (dof_number[row_local]) */
  const int32_t* row_local; /* [nb_row] local dof of each row (to scatter the solution back) */
  const int32_t* cols; /* [nb_value] global column ids */
  const double* values; /* [nb_value] */
  const double* rhs; /* [nb_row] right-hand side of the own rows */
} b2f_ij_view;
/* Own-row extraction + global numbering on the device (the reference loops over own dofs and entries on the host,
 * HypreDoFLinearSystem.cc:474-519, then copies four arrays to the device, :587-599).  dof_number may be NULL (local ids
 * are kept: single-rank runs).  Call after the Dirichlet treatment; synchronises `stream` once (two totals). */
int b2f_to_hypre_ij(b2f_handle h, const int32_t* dof_number, b2f_ij_view* out, b2f_stream stream);
int b2f_get_system(b2f_handle h, b2f_system_view* out);
/* Device pointers of the gather plan built by b2f_compute_sparsity (read-only; for inspection and tests). */
int b2f_get_plan(b2f_handle h, b2f_plan_view* out);
int b2f_get_stats(b2f_handle h, b2f_stats* out);

/* BSRMatrix::setValue / getValue by scalar (row dof, col dof), BSRFormat.h:176-189 (host convenience,
 * synchronous).  B2F_ERR_NOT_FOUND when the block is not in the pattern. */
int b2f_set_value(b2f_handle h, int64_t row_dof, int64_t col_dof, double value);
int b2f_get_value(b2f_handle h, int64_t row_dof, int64_t col_dof, double* value);
/* BSRMatrix::dump (BSRFormat.h:285-307), same text format. */
int b2f_dump(b2f_handle h, const char* filename);

/* DoFLinearSystem::clearValues (HypreDoFLinearSystem.cc:184-191) + rhs.fill(0). */
int b2f_clear_values(b2f_handle h, b2f_stream stream);
int b2f_rhs_fill(b2f_handle h, double value, b2f_stream stream);
/* BoundaryConditions{2D,3D}::applyConstantSourceToRhs (ArcaneFemFunctionsGpu.h:402-429,605,695). */
int b2f_rhs_constant_source(b2f_handle h, double qdot, b2f_stream stream);
/* Source term of the transient heat module (modules/heat/FemModule.cc:480-495): rhs[dof(n,0)] += (node_field[n] * scale) * |cell| / Nv
 * over the cells around every own node that node_skip (m_node_is_temperature_fixed; may be NULL) does not exclude;
 * node_field = last step's temperature, scale = 1/dt.  Triangles use the module's signed 2D area (:605-612). */
int b2f_rhs_nodal_source(b2f_handle h, const double* node_field, double scale, const uint8_t* node_skip, b2f_stream stream);
/* BoundaryConditions{2D,3D}::applyNeumannToRhs (ArcaneFemFunctionsGpu.h:621-677,711-768). */
int b2f_rhs_neumann(b2f_handle h, const b2f_neumann* bc, b2f_stream stream);
/* Convection (Robin) boundary term of the heat module, modules/heat/FemModule.cc:628-672 (_computeElementMatrixEDGE2) scattered
 * by _assembleBilinearOperatorEDGE2 (:813-840): for every boundary face, K[a,b] += coefficient * |face| / (Nvf (Nvf+1)) * (1 + delta_ab)
 * over the face's nodes, own rows only - edges of a 2D mesh: h L/6 [[2,1],[1,2]]; triangles of a 3D mesh (twin, not in the
 * reference): h A/12 (1 + delta).  face_node: [nb_face * dim] device node ids.  Added to the current values (call after
 * b2f_assemble); block_size 1.  b2f_stats.columns_not_found counts face edges that are not in the pattern.  The matching
 * right-hand side h * Text * |face| / Nvf is b2f_rhs_neumann with value = h * Text (modules/heat/FemModule.cc:432-470). */
int b2f_add_face_mass(b2f_handle h, int64_t nb_face, const int32_t* face_node, double coefficient, b2f_stream stream);
/* applyDirichletToNodeGroupViaPenalty (ArcaneFemFunctionsGpu.h:434-456). */
int b2f_dirichlet_flag_penalty(b2f_handle h, const int32_t* node_ids, int64_t n, int32_t dof_index, double value, double penalty, b2f_stream stream);
/* WeakPenalty (modules/testlab/FemModule.cc:687-717; CPU-only in the reference): a_ii += P at b2f_dirichlet_apply, rhs_i = P*g. */
int b2f_dirichlet_flag_weak_penalty(b2f_handle h, const int32_t* node_ids, int64_t n, int32_t dof_index, double value, double penalty, b2f_stream stream);
/* applyDirichletToNodeGroupViaRowElimination<MODE> (ArcaneFemFunctionsGpu.h:461-482). */
int b2f_dirichlet_flag_elimination(b2f_handle h, const int32_t* node_ids, int64_t n, int32_t dof_index, double value, int32_t mode, b2f_stream stream);
/* _applyRowElimination + _applyForcedValuesToLhs (HypreDoFLinearSystem.cc:319-382), plus row+column
 * elimination with the semantics of AlephDoFLinearSystem.cc:501-583, on the device-resident matrix. */
int b2f_dirichlet_apply(b2f_handle h, b2f_stream stream);

/* Device BSR SpMV check: y = A x, x and y of length BS*nb_node. */
int b2f_spmv(b2f_handle h, const double* x, double* y, b2f_stream stream);

/*
 * Host-buffer entry point (what a caller without device-resident data uses; timed as `e2e` by bench.py):
 * copies coordinates host->device into the handle's own staging buffer, assembles, and copies the values
 * device->host.  Pinned host memory recommended.  values_host may be NULL (skip the read-back).
 */
int b2f_assemble_host(b2f_handle h, const b2f_operator* op, const double* coords_host, double* values_host, b2f_stream stream);

/*
 * Multi-GPU (one rank per GPU, Arcane-style decomposition: the local mesh holds the rank's own cells plus the ghost-cell
 * layer, b2f_mesh_view::cell_is_own marks the former).  Only own cells are assembled, into all local rows; rows of
 * ghost nodes then hold this rank's partial sums.  b2f_ghost_exchange sends them to the owner rank over NCCL
 * (grouped ncclSend/ncclRecv per neighbour), adds them into the owner's rows and zeroes the ghost rows, so owner rows
 * equal the reference's (which computes the ghost layer redundantly instead, femutils/BSRFormat.h:807-815).
 * `comm` is an ncclComm_t.  Plan inputs are host arrays over the local nodes: owner rank and global id
 * (Arcane: Node::owner(), uniqueId()).  mesh.node_is_own must be NULL in this mode.
 */
int b2f_ghost_plan(b2f_handle h, void* nccl_comm, int32_t rank, int32_t world, const int32_t* node_owner_rank_host, const int64_t* node_global_id_host,
                   b2f_stream stream);
int b2f_ghost_exchange(b2f_handle h, void* nccl_comm, b2f_stream stream);
/* One multi-GPU assembly step, b2f_assemble followed by b2f_ghost_exchange with the exchange hidden: the ghost rows are
 * assembled first and travel on a high-priority stream of the handle while the remaining rows are assembled on `stream`;
 * the received sums are added at the end, on `stream`.  Same values as the two calls in sequence (every row is written by
 * one launch, the additions happen in the same order).  Falls back to the sequence when the operator accumulates, the
 * default gather kernel is not in use, or the sent rows span more than a quarter of the local numbering. */
int b2f_assemble_exchange(b2f_handle h, const b2f_operator* op, void* nccl_comm, b2f_stream stream);
/* Host-buffer entry point of the multi-GPU step (b2f_assemble_host + b2f_assemble_exchange in one pipeline; what bench.py times as
 * `e2e` at N > 1): coordinates go up as a growing prefix, row chunks are assembled as they land, the ghost rows travel as soon as
 * the last of them is assembled, and every chunk's values go back to the host while the next chunk is assembled - chunks whose
 * rows travel or receive partial sums after the sums have been added.  Same values as b2f_assemble_exchange followed by a copy.
 * Pinned host memory recommended; values_host may be NULL.  Synchronises before returning.  Falls back to copy / step / copy
 * when the default gather kernel is not in use, the operator accumulates, or the mesh has fewer than 65536 nodes. */
int b2f_assemble_exchange_host(b2f_handle h, const b2f_operator* op, void* nccl_comm, const double* coords_host, double* values_host, b2f_stream stream);
/* NCCL bootstrap helpers so a host without its own NCCL binding can create the communicator. */
int b2f_nccl_unique_id(void* id128_out_host);
int b2f_nccl_comm_init(const void* id128_host, int32_t rank, int32_t world, void** comm_out);
int b2f_nccl_comm_destroy(void* comm);

/*
 * Allocation-free own-row extraction.  NO COUNTERPART IN THE REFERENCE: added for
 * callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Same result as b2f_to_hypre_ij (own rows of a CSR matrix in local order, global row / column ids, bit-identical
 * values), but handle-free: the caller passes every input and every output as device pointers, and the two scan arrays
 * the extraction needs (row slot and value offset per row, plus per-tile sums) live in a file-scope
 * `static __device__` array of the library instead of the handle's workspace.  The entry point calls no cudaMalloc,
 * cudaFree or allocating library routine and does not synchronise.
 *   limit   : nb_row <= B2F_NO_ALLOC_MAX_ROWS scalar rows (the array is 8 MiB + 12 KiB, resident on every GPU that loads
 *             the library until the process exits); larger sizes are refused with B2F_ERR_INVALID_ARGUMENT.
 *   device  : the calling thread's current device (cudaSetDevice); all pointers must belong to it.  The address of the
 *             array is looked up once per device on the first call there - a call with nb_row = 0 does only that, for
 *             callers that want the set-up outside a capture.
 *   sharing : every call on a device uses the same array.  Nothing in it needs clearing (each call writes every word
 *             it later reads), but two calls that overlap on the GPU (different streams, no ordering) race; calls on
 *             one stream, or ordered by events, are safe.
 *   inputs  : row_ptr [nb_row] (or nb_row + 1) first entry of each row; row_ncols [nb_row] row lengths, NULL = row_ptr[r+1]-row_ptr[r];
 *             cols / values the CSR arrays; rhs, node_is_own (per block row: row r belongs to node r / block_size) and
 *             dof_number (global numbering; NULL keeps local ids) may be NULL.
 *   outputs : arrays of the capacities given; rows or entries beyond a capacity are dropped, never written.
 *             totals_out (device, 2 x int64: own rows, their entries) lets the caller check them against the capacities.
 */
#define B2F_NO_ALLOC_MAX_ROWS (1 << 20)
typedef struct b2f_ij_buffers {
  int64_t capacity_rows; /* length of ncols, rows, row_local, rhs */
  int64_t capacity_values; /* length of cols, values */
  int32_t* ncols;
  int32_t* rows;
  int32_t* row_local;
  int32_t* cols;
  double* values;
  double* rhs; /* may be NULL */
  int64_t* totals_out; /* [2] */
} b2f_ij_buffers;
int b2f_ij_own_rows_no_alloc(const int32_t* row_ptr, const int32_t* row_ncols, const int32_t* cols, const double* values, const double* rhs,
                             const uint8_t* node_is_own, const int32_t* dof_number, int64_t nb_row, int32_t block_size, const b2f_ij_buffers* out,
                             b2f_stream stream);

#ifdef __cplusplus
}
#endif
#endif /* B2F_H */
