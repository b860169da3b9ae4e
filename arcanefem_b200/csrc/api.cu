// api.cu — the extern "C" surface declared in include/b2f.h.
#include <cstring>
#include <fstream>
#include "internal.h"
#include "b2f/assemble_args.h"
#include <algorithm>

namespace b2f {

static thread_local char g_error[1024] = "";

void set_error(const char* fmt, ...)
{
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_error, sizeof(g_error), fmt, ap);
  va_end(ap);
}

int fail(int status, const char* fmt, ...)
{
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_error, sizeof(g_error), fmt, ap);
  va_end(ap);
  return status;
}

int fail_cuda(cudaError_t e, const char* what)
{
  snprintf(g_error, sizeof(g_error), "CUDA error %d (%s) in %s", (int)e, cudaGetErrorString(e), what);
  return B2F_ERR_CUDA;
}

} // namespace b2f

using namespace b2f;

#define B2F_CHECK_HANDLE(h)                                        \
  if ((h) == nullptr)                                              \
  return fail(B2F_ERR_INVALID_ARGUMENT, "%s: null handle", __func__)
#define B2F_NEED_PATTERN(h)                                        \
  if (!(h)->has_pattern)                                           \
  return fail(B2F_ERR_STATE, "%s: b2f_compute_sparsity has not been called", __func__)
// every entry that launches, copies or allocates runs on the device the handle was created on (a single-process caller
// with several devices may have another one current)
#define B2F_ON_DEVICE(h) B2F_CUDA(cudaSetDevice((h)->device))

#ifndef B2F_HOST_PIPELINE_CHUNKS
#define B2F_HOST_PIPELINE_CHUNKS 8
#endif

extern "C" {

int b2f_version(void) { return B2F_VERSION_MAJOR * 1000 + B2F_VERSION_MINOR; }

const char* b2f_last_error_string(void) { return g_error; }

const char* b2f_status_string(int status)
{
  switch (status) {
  case B2F_OK: return "B2F_OK";
  case B2F_ERR_INVALID_ARGUMENT: return "B2F_ERR_INVALID_ARGUMENT";
  case B2F_ERR_NOT_IMPLEMENTED: return "B2F_ERR_NOT_IMPLEMENTED";
  case B2F_ERR_CUDA: return "B2F_ERR_CUDA";
  case B2F_ERR_STATE: return "B2F_ERR_STATE";
  case B2F_ERR_NOT_FOUND: return "B2F_ERR_NOT_FOUND";
  case B2F_ERR_OVERFLOW: return "B2F_ERR_OVERFLOW";
  case B2F_ERR_NCCL: return "B2F_ERR_NCCL";
  }
  return "B2F_ERR_UNKNOWN";
}

int b2f_create(const b2f_mesh_view* mesh, const b2f_options* options, b2f_handle* out)
{
  if (!mesh || !options || !out)
    return fail(B2F_ERR_INVALID_ARGUMENT, "b2f_create: null argument");
  *out = nullptr;
  // BSRFormat::initialize: "Only supports 2D and 3D" (femutils/BSRFormat.h:393-394)
  if (mesh->dim != 2 && mesh->dim != 3)
    return fail(B2F_ERR_NOT_IMPLEMENTED, "BSRFormat(initialize): Only supports 2D and 3D");
  if (!((mesh->dim == 2 && mesh->nodes_per_cell == 3) || (mesh->dim == 3 && mesh->nodes_per_cell == 4)))
    return fail(B2F_ERR_NOT_IMPLEMENTED, "only P1 triangles (dim 2) and P1 tetrahedra (dim 3) are supported (dim=%d, nodes_per_cell=%d)", mesh->dim, mesh->nodes_per_cell);
  // BSRMatrix::initialize argument checks (femutils/BSRFormat.h:121-125)
  if (options->block_size <= 0 || mesh->nb_node <= 0 || mesh->nb_cell < 0)
    return fail(B2F_ERR_INVALID_ARGUMENT, "BSRMatrix(initialize): arguments should be positive and not null (block_size=%d, nb_cell=%lld and nb_row=%lld)",
                options->block_size, (long long)mesh->nb_cell, (long long)mesh->nb_node);
  if (options->block_size > 3)
    return fail(B2F_ERR_NOT_IMPLEMENTED, "block_size %d: only 1, 2 and 3 dofs per node are built", options->block_size);
  if (options->block_size > mesh->nb_node)
    return fail(B2F_ERR_INVALID_ARGUMENT, "BSRMatrix(initialize): block_size should be less than nb_row");
  if (mesh->nb_node > 2147483647ll)
    return fail(B2F_ERR_OVERFLOW, "nb_node=%lld does not fit Int32 node ids", (long long)mesh->nb_node);
  if (!mesh->cell_node && mesh->nb_cell > 0)
    return fail(B2F_ERR_INVALID_ARGUMENT, "b2f_create: cell_node is null");
  // the kernels read a tetrahedron's four node ids with one 16-byte load and coordinates with 8-byte loads: a misaligned
  // view (a sliced tensor, an offset connectivity span) would fault on the device and take the CUDA context with it
  if (mesh->nodes_per_cell == 4 && (reinterpret_cast<uintptr_t>(mesh->cell_node) & 15u) != 0)
    return fail(B2F_ERR_INVALID_ARGUMENT, "b2f_create: cell_node of a tetrahedral mesh must be 16-byte aligned (got %p)", (const void*)mesh->cell_node);
  if ((reinterpret_cast<uintptr_t>(mesh->cell_node) & 3u) != 0 || (reinterpret_cast<uintptr_t>(mesh->coords) & 7u) != 0)
    return fail(B2F_ERR_INVALID_ARGUMENT, "b2f_create: cell_node must be 4-byte and coords 8-byte aligned");
  if (options->assembly_variant < B2F_ASM_DEFAULT || options->assembly_variant > B2F_ASM_NODE_GATHER_SLOTS)
    return fail(B2F_ERR_INVALID_ARGUMENT, "b2f_create: unknown assembly_variant %d", options->assembly_variant);
  if (options->sparsity_build < B2F_SPARSITY_AUTO || options->sparsity_build > B2F_SPARSITY_LISTS)
    return fail(B2F_ERR_INVALID_ARGUMENT, "b2f_create: unknown sparsity_build %d", options->sparsity_build);
  if (options->pattern_layout != B2F_PATTERN_DIAG_FIRST && options->pattern_layout != B2F_PATTERN_DIAG_LAST)
    return fail(B2F_ERR_INVALID_ARGUMENT, "b2f_create: unknown pattern_layout %d", options->pattern_layout);
  if (options->pattern_layout == B2F_PATTERN_DIAG_LAST && options->assembly_variant == B2F_ASM_NODE_GATHER_SLOTS)
    return fail(B2F_ERR_NOT_IMPLEMENTED, "b2f_create: the first-generation slot-word kernel (kept for A/B measurements) only knows the diagonal-first layout");
  if (options->reorder < B2F_REORDER_NONE || options->reorder > B2F_REORDER_PENCIL)
    return fail(B2F_ERR_INVALID_ARGUMENT, "b2f_create: unknown reorder %d", options->reorder);
  if (options->fan_mode < B2F_FAN_AUTO || options->fan_mode > B2F_FAN_LOCKSTEP)
    return fail(B2F_ERR_INVALID_ARGUMENT, "b2f_create: unknown fan_mode %d", options->fan_mode);
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0)
    return fail(B2F_ERR_CUDA, "no CUDA device available (%s); libb200fem has no CPU fallback", e != cudaSuccess ? cudaGetErrorString(e) : "device count 0");
  b2f_context_s* h = new b2f_context_s();
  h->mesh = *mesh;
  h->opt = *options;
  h->bs = options->block_size;
  h->nv = mesh->nodes_per_cell;
  h->per_block = options->use_csr == 0; // order_values_per_block = !does_linear_system_use_csr, BSRFormat.h:405
  h->variant = options->assembly_variant == B2F_ASM_DEFAULT ? B2F_ASM_NODE_GATHER : options->assembly_variant;
  h->diag_last = options->pattern_layout == B2F_PATTERN_DIAG_LAST ? 1 : 0;
  cudaGetDevice(&h->device);
  *out = h;
  return B2F_OK;
}

int b2f_destroy(b2f_handle h)
{
  if (!h)
    return B2F_OK;
  DevBuf* bufs[] = { &h->row_index, &h->columns, &h->nb_nz_per_row, &h->values, &h->nc_ptr, &h->nc_list, &h->nc_slots, &h->fan_ptr, &h->fan_steps, &h->fan_cell, &h->csr_rows, &h->csr_nbcol, &h->csr_cols, &h->rhs,
                     &h->forced_info, &h->forced_value, &h->elim_info, &h->elim_value, &h->counters, &h->ws_keys_a, &h->ws_keys_b, &h->ws_tmp, &h->ws_i32a, &h->ws_i64,
                     &h->coords_stage, &h->ghost.send_index, &h->ghost.recv_index, &h->ghost.send_vals, &h->ghost.recv_vals, &h->ghost.recv_dst, &h->ghost.recv_src,
                     &h->ij_flag, &h->ij_off, &h->ij_ncols, &h->ij_rows, &h->ij_row_local, &h->ij_cols, &h->ij_values, &h->ij_rhs,
                     &h->perm_rows, &h->r_row_index, &h->r_columns, &h->r_coords, &h->r_slot_map };
  for (DevBuf* b : bufs)
    b->release();
  for (cudaEvent_t e : h->pipe_events)
    cudaEventDestroy(e);
  if (h->pipe_h2d)
    cudaStreamDestroy(h->pipe_h2d);
  if (h->pipe_d2h)
    cudaStreamDestroy(h->pipe_d2h);
  for (cudaEvent_t e : h->comm_events)
    if (e)
      cudaEventDestroy(e);
  if (h->comm_stream)
    cudaStreamDestroy(h->comm_stream);
  delete h;
  return B2F_OK;
}

int b2f_set_coords(b2f_handle h, const double* coords)
{
  B2F_CHECK_HANDLE(h);
  if ((reinterpret_cast<uintptr_t>(coords) & 7u) != 0)
    return fail(B2F_ERR_INVALID_ARGUMENT, "b2f_set_coords: coords must be 8-byte aligned");
  h->mesh.coords = coords;
  return B2F_OK;
}

int b2f_compute_sparsity(b2f_handle h, b2f_stream stream)
{
  B2F_CHECK_HANDLE(h);
  B2F_ON_DEVICE(h);
  return build_pattern(h, (cudaStream_t)stream);
}

int b2f_release_workspace(b2f_handle h)
{
  B2F_CHECK_HANDLE(h);
  B2F_ON_DEVICE(h);
  h->ws_keys_a.release();
  h->ws_keys_b.release();
  h->ws_tmp.release();
  h->ws_i32a.release();
  h->ws_i64.release();
  return B2F_OK;
}

int b2f_assemble(b2f_handle h, const b2f_operator* op, b2f_stream stream)
{
  B2F_CHECK_HANDLE(h);
  B2F_NEED_PATTERN(h);
  B2F_ON_DEVICE(h);
  if (!op)
    return fail(B2F_ERR_INVALID_ARGUMENT, "b2f_assemble: null operator");
  if (!h->mesh.coords)
    return fail(B2F_ERR_INVALID_ARGUMENT, "b2f_assemble: mesh coordinates are null");
  return assemble(h, op, (cudaStream_t)stream);
}

int b2f_assemble_with_source(b2f_handle h, const b2f_operator* op, double qdot, b2f_stream stream)
{
  B2F_CHECK_HANDLE(h);
  B2F_NEED_PATTERN(h);
  B2F_ON_DEVICE(h);
  if (!op)
    return fail(B2F_ERR_INVALID_ARGUMENT, "b2f_assemble_with_source: null operator");
  if (!h->mesh.coords)
    return fail(B2F_ERR_INVALID_ARGUMENT, "b2f_assemble_with_source: mesh coordinates are null");
  return assemble(h, op, (cudaStream_t)stream, &qdot);
}

int b2f_user_assemble_begin(b2f_handle h, int32_t nodes_per_cell, int32_t block_size, void* args_out, size_t args_size, int32_t* variant_out, int32_t* per_block_out,
                            int64_t* nb_values_out, b2f_stream stream)
{
  B2F_CHECK_HANDLE(h);
  B2F_NEED_PATTERN(h);
  B2F_ON_DEVICE(h);
  if (!args_out || !variant_out || !per_block_out || !nb_values_out)
    return fail(B2F_ERR_INVALID_ARGUMENT, "b2f_user_assemble_begin: null output");
  if (args_size != sizeof(AssembleArgs))
    return fail(B2F_ERR_INVALID_ARGUMENT, "b2f_user_assemble_begin: caller was compiled against another b2f::AssembleArgs (%zu bytes, library has %zu)",
                args_size, sizeof(AssembleArgs));
  if (nodes_per_cell != h->nv || block_size != h->bs)
    return fail(B2F_ERR_INVALID_ARGUMENT, "b2f_user_assemble_begin: operator is for %d nodes per cell, block size %d; the handle has %d and %d", nodes_per_cell,
                block_size, h->nv, h->bs);
  if (!h->mesh.coords)
    return fail(B2F_ERR_INVALID_ARGUMENT, "b2f_user_assemble_begin: mesh coordinates are null");
  B2F_CUDA(cudaMemsetAsync(h->counters.p, 0, sizeof(unsigned long long) * 2, (cudaStream_t)stream));
  if (h->reordered) { // a caller-compiled operator that runs on the fan kernel reads the internal coordinates too
    const int rc = permute_coords(h, (cudaStream_t)stream);
    if (rc != B2F_OK)
      return rc;
  }
  fill_assemble_args(h, static_cast<AssembleArgs*>(args_out));
  *variant_out = h->variant == B2F_ASM_DEFAULT ? B2F_ASM_NODE_GATHER : h->variant;
  *per_block_out = h->per_block ? 1 : 0;
  *nb_values_out = h->nb_nz;
  return B2F_OK;
}

int b2f_reset_values(b2f_handle h, b2f_stream stream)
{
  B2F_CHECK_HANDLE(h);
  B2F_NEED_PATTERN(h);
  B2F_ON_DEVICE(h);
  B2F_CUDA(cudaMemsetAsync(h->values.p, 0, sizeof(double) * (size_t)h->nb_nz, (cudaStream_t)stream));
  return B2F_OK;
}

int b2f_get_bsr(b2f_handle h, b2f_bsr_view* out)
{
  B2F_CHECK_HANDLE(h);
  B2F_NEED_PATTERN(h);
  B2F_ON_DEVICE(h);
  if (!out)
    return fail(B2F_ERR_INVALID_ARGUMENT, "b2f_get_bsr: null output");
  out->block_size = h->bs;
  out->layout = h->per_block ? B2F_LAYOUT_PER_BLOCK : B2F_LAYOUT_PER_ROW;
  out->nb_row = h->mesh.nb_node;
  out->nb_col = h->nb_col;
  out->nb_nz = h->nb_nz;
  out->nb_edge = h->nb_edge;
  out->row_index = h->row_index.as<int32_t>();
  out->columns = h->columns.as<int32_t>();
  out->nb_nz_per_row = h->nb_nz_per_row.as<int32_t>();
  out->values = h->values.as<double>();
  return B2F_OK;
}

int b2f_to_csr(b2f_handle h, b2f_csr_view* out, b2f_stream stream)
{
  B2F_CHECK_HANDLE(h);
  B2F_NEED_PATTERN(h);
  B2F_ON_DEVICE(h);
  if (!out)
    return fail(B2F_ERR_INVALID_ARGUMENT, "b2f_to_csr: null output");
  if (!h->has_csr) {
    int rc = to_csr(h, (cudaStream_t)stream);
    if (rc != B2F_OK)
      return rc;
  }
  out->nb_row = h->mesh.nb_node * h->bs;
  out->nb_value = h->nb_nz;
  if (h->bs == 1) {
    out->rows = h->row_index.as<int32_t>();
    out->rows_nb_column = h->nb_nz_per_row.as<int32_t>();
    out->columns = h->columns.as<int32_t>();
  }
  else {
    out->rows = h->csr_rows.as<int32_t>();
    out->rows_nb_column = h->csr_nbcol.as<int32_t>();
    out->columns = h->csr_cols.as<int32_t>();
  }
  out->values = h->values.as<double>();
  return B2F_OK;
}

int b2f_csr_global_numbering(b2f_handle h, const int32_t* dof_number, int32_t* rows_global_out, int32_t* cols_global_out, b2f_stream stream)
{
  B2F_CHECK_HANDLE(h);
  B2F_NEED_PATTERN(h);
  B2F_ON_DEVICE(h);
  if (!dof_number)
    return fail(B2F_ERR_INVALID_ARGUMENT, "b2f_csr_global_numbering: null numbering");
  return csr_global_numbering(h, dof_number, rows_global_out, cols_global_out, (cudaStream_t)stream);
}

int b2f_to_hypre_ij(b2f_handle h, const int32_t* dof_number, b2f_ij_view* out, b2f_stream stream)
{
  B2F_CHECK_HANDLE(h);
  B2F_NEED_PATTERN(h);
  B2F_ON_DEVICE(h);
  if (!out)
    return fail(B2F_ERR_INVALID_ARGUMENT, "b2f_to_hypre_ij: null output");
  return to_hypre_ij(h, dof_number, out, (cudaStream_t)stream);
}

int b2f_get_system(b2f_handle h, b2f_system_view* out)
{
  B2F_CHECK_HANDLE(h);
  B2F_NEED_PATTERN(h);
  B2F_ON_DEVICE(h);
  if (!out)
    return fail(B2F_ERR_INVALID_ARGUMENT, "b2f_get_system: null output");
  out->nb_dof = h->nb_dof;
  out->rhs = h->rhs.as<double>();
  out->forced_info = h->forced_info.as<uint8_t>();
  out->forced_value = h->forced_value.as<double>();
  out->elimination_info = h->elim_info.as<uint8_t>();
  out->elimination_value = h->elim_value.as<double>();
  return B2F_OK;
}

int b2f_get_plan(b2f_handle h, b2f_plan_view* out)
{
  B2F_CHECK_HANDLE(h);
  B2F_NEED_PATTERN(h);
  B2F_ON_DEVICE(h);
  if (!out)
    return fail(B2F_ERR_INVALID_ARGUMENT, "b2f_get_plan: null output");
  out->nb_node = h->mesh.nb_node;
  out->nb_pair = h->nb_pair;
  out->nb_fan_step = h->fan_ok ? h->nb_fan_step : 0;
  out->nc_ptr = h->nc_ptr.as<int32_t>();
  out->nc_list = h->nc_list.as<uint32_t>();
  out->fan_ptr = h->fan_ok ? h->fan_ptr.as<int32_t>() : nullptr;
  out->fan_steps = h->fan_ok ? h->fan_steps.as<uint32_t>() : nullptr;
  out->fan_cell = h->fan_ok ? h->fan_cell.as<uint32_t>() : nullptr;
  return B2F_OK;
}

int b2f_get_stats(b2f_handle h, b2f_stats* out)
{
  B2F_CHECK_HANDLE(h);
  if (!out)
    return fail(B2F_ERR_INVALID_ARGUMENT, "b2f_get_stats: null output");
  B2F_ON_DEVICE(h);
  memset(out, 0, sizeof(*out));
  out->nb_edge = h->nb_edge;
  out->nb_col = h->nb_col;
  out->nb_nz = h->nb_nz;
  out->fan_steps = h->nb_fan_step;
  out->sparsity_by_lists = h->pattern_by_lists ? 1 : 0;
  out->fan_lockstep = h->fan_lockstep ? 1 : 0;
  out->fan_lockstep_bodies_milli = (int64_t)(h->fan_lockstep_bodies * 1000.0 + 0.5);
  out->reordered = h->reordered ? 1 : 0;
  if (h->counters.p) {
    unsigned long long c[4] = { 0, 0, 0, 0 };
    B2F_CUDA(cudaMemcpy(c, h->counters.p, sizeof(c), cudaMemcpyDeviceToHost));
    out->columns_not_found = (int64_t)c[0];
    out->gather_fallback_tiles = (int64_t)c[1];
  }
  out->workspace_bytes = (int64_t)(h->ws_keys_a.bytes + h->ws_keys_b.bytes + h->ws_tmp.bytes + h->ws_i32a.bytes + h->ws_i64.bytes);
  const DevBuf* res[] = { &h->row_index, &h->columns, &h->nb_nz_per_row, &h->values, &h->nc_ptr, &h->nc_list, &h->nc_slots, &h->fan_ptr, &h->fan_steps, &h->fan_cell, &h->csr_rows, &h->csr_nbcol, &h->csr_cols, &h->rhs,
                          &h->forced_info, &h->forced_value, &h->elim_info, &h->elim_value, &h->counters, &h->coords_stage, &h->perm_rows, &h->r_row_index, &h->r_columns,
                          &h->r_coords, &h->r_slot_map };
  for (const DevBuf* b : res)
    out->resident_bytes += (int64_t)b->bytes;
  return B2F_OK;
}

// findValueIndex, femutils/BSRFormat.h:145-171 (host walk over one row copied from the device).
static int find_value_index(b2f_handle h, int64_t row, int64_t col, int64_t* idx)
{
  const int bs = h->bs;
  if (row < 0 || col < 0 || row >= h->nb_dof || col >= h->nb_dof)
    return fail(B2F_ERR_INVALID_ARGUMENT, "dof (%lld,%lld) out of range", (long long)row, (long long)col);
  const int64_t br = row / bs, bc = col / bs;
  int32_t rr[2];
  B2F_CUDA(cudaMemcpy(rr, h->row_index.as<int32_t>() + br, sizeof(rr), cudaMemcpyDeviceToHost));
  const int k = rr[1] - rr[0];
  std::vector<int32_t> cols((size_t)k);
  B2F_CUDA(cudaMemcpy(cols.data(), h->columns.as<int32_t>() + rr[0], sizeof(int32_t) * (size_t)k, cudaMemcpyDeviceToHost));
  for (int x = 0; x < k; ++x) {
    if (cols[x] == bc) {
      const int i = (int)(row % bs), j = (int)(col % bs);
      if (!h->per_block)
        *idx = (int64_t)rr[0] * bs * bs + (int64_t)bs * x + (int64_t)i * bs * k + j;
      else
        *idx = ((int64_t)rr[0] + x) * bs * bs + (i * bs + j);
      return B2F_OK;
    }
  }
  return fail(B2F_ERR_NOT_FOUND, "BSRMatrix(findValueIndex): Value not found");
}

int b2f_set_value(b2f_handle h, int64_t row_dof, int64_t col_dof, double value)
{
  B2F_CHECK_HANDLE(h);
  B2F_NEED_PATTERN(h);
  B2F_ON_DEVICE(h);
  int64_t idx = 0;
  int rc = find_value_index(h, row_dof, col_dof, &idx);
  if (rc != B2F_OK)
    return rc;
  B2F_CUDA(cudaMemcpy(h->values.as<double>() + idx, &value, sizeof(double), cudaMemcpyHostToDevice));
  return B2F_OK;
}

int b2f_get_value(b2f_handle h, int64_t row_dof, int64_t col_dof, double* value)
{
  B2F_CHECK_HANDLE(h);
  B2F_NEED_PATTERN(h);
  B2F_ON_DEVICE(h);
  if (!value)
    return fail(B2F_ERR_INVALID_ARGUMENT, "b2f_get_value: null output");
  int64_t idx = 0;
  int rc = find_value_index(h, row_dof, col_dof, &idx);
  if (rc != B2F_OK)
    return rc;
  B2F_CUDA(cudaMemcpy(value, h->values.as<double>() + idx, sizeof(double), cudaMemcpyDeviceToHost));
  return B2F_OK;
}

// BSRMatrix::dump, femutils/BSRFormat.h:285-307 (same text layout).
int b2f_dump(b2f_handle h, const char* filename)
{
  B2F_CHECK_HANDLE(h);
  B2F_NEED_PATTERN(h);
  B2F_ON_DEVICE(h);
  if (!filename)
    return fail(B2F_ERR_INVALID_ARGUMENT, "b2f_dump: null filename");
  const int64_t nn = h->mesh.nb_node;
  std::vector<int32_t> ri((size_t)nn + 1), co((size_t)h->nb_col);
  std::vector<double> va((size_t)h->nb_nz);
  B2F_CUDA(cudaMemcpy(ri.data(), h->row_index.p, sizeof(int32_t) * ri.size(), cudaMemcpyDeviceToHost));
  B2F_CUDA(cudaMemcpy(co.data(), h->columns.p, sizeof(int32_t) * co.size(), cudaMemcpyDeviceToHost));
  B2F_CUDA(cudaMemcpy(va.data(), h->values.p, sizeof(double) * va.size(), cudaMemcpyDeviceToHost));
  std::ofstream file(filename);
  if (!file)
    return fail(B2F_ERR_INVALID_ARGUMENT, "b2f_dump: cannot open %s", filename);
  file << "size :" << h->nb_nz << "\n";
  for (int64_t i = 0; i < nn; ++i) {
    file << ri[i] << " ";
    for (int32_t j = ri[i] + 1; j < ri[i + 1]; ++j)
      file << "  ";
  }
  file << "\n";
  for (int64_t i = 0; i < h->nb_col; ++i)
    file << co[i] << " ";
  file << "\n";
  for (int64_t i = 0; i < h->nb_nz; ++i)
    file << va[i] << " ";
  file << "\n";
  return B2F_OK;
}

int b2f_clear_values(b2f_handle h, b2f_stream stream)
{
  B2F_CHECK_HANDLE(h);
  B2F_NEED_PATTERN(h);
  B2F_ON_DEVICE(h);
  return clear_values(h, (cudaStream_t)stream);
}

int b2f_rhs_fill(b2f_handle h, double value, b2f_stream stream)
{
  B2F_CHECK_HANDLE(h);
  B2F_NEED_PATTERN(h);
  B2F_ON_DEVICE(h);
  return rhs_fill(h, value, (cudaStream_t)stream);
}

int b2f_rhs_constant_source(b2f_handle h, double qdot, b2f_stream stream)
{
  B2F_CHECK_HANDLE(h);
  B2F_NEED_PATTERN(h);
  B2F_ON_DEVICE(h);
  if (!h->mesh.coords)
    return fail(B2F_ERR_INVALID_ARGUMENT, "b2f_rhs_constant_source: mesh coordinates are null");
  return rhs_constant_source(h, qdot, (cudaStream_t)stream);
}

int b2f_rhs_nodal_source(b2f_handle h, const double* node_field, double scale, const uint8_t* node_skip, b2f_stream stream)
{
  B2F_CHECK_HANDLE(h);
  B2F_NEED_PATTERN(h);
  B2F_ON_DEVICE(h);
  if (!node_field)
    return fail(B2F_ERR_INVALID_ARGUMENT, "b2f_rhs_nodal_source: null node field");
  if (!h->mesh.coords)
    return fail(B2F_ERR_INVALID_ARGUMENT, "b2f_rhs_nodal_source: mesh coordinates are null");
  return rhs_nodal_source(h, node_field, scale, node_skip, (cudaStream_t)stream);
}

int b2f_rhs_neumann(b2f_handle h, const b2f_neumann* bc, b2f_stream stream)
{
  B2F_CHECK_HANDLE(h);
  B2F_NEED_PATTERN(h);
  B2F_ON_DEVICE(h);
  if (!bc || (bc->nb_face > 0 && !bc->face_node))
    return fail(B2F_ERR_INVALID_ARGUMENT, "b2f_rhs_neumann: null boundary condition / face list");
  return rhs_neumann(h, bc, (cudaStream_t)stream);
}

int b2f_add_face_mass(b2f_handle h, int64_t nb_face, const int32_t* face_node, double coefficient, b2f_stream stream)
{
  B2F_CHECK_HANDLE(h);
  B2F_NEED_PATTERN(h);
  B2F_ON_DEVICE(h);
  if (h->bs != 1)
    return fail(B2F_ERR_INVALID_ARGUMENT, "b2f_add_face_mass needs block_size 1 (got %d)", h->bs);
  if (nb_face < 0 || (nb_face > 0 && !face_node))
    return fail(B2F_ERR_INVALID_ARGUMENT, "b2f_add_face_mass: bad face list");
  if (!h->mesh.coords)
    return fail(B2F_ERR_INVALID_ARGUMENT, "b2f_add_face_mass: mesh coordinates are null");
  return add_face_mass(h, nb_face, face_node, coefficient, (cudaStream_t)stream);
}

int b2f_dirichlet_flag_penalty(b2f_handle h, const int32_t* node_ids, int64_t n, int32_t dof_index, double value, double penalty, b2f_stream stream)
{
  B2F_CHECK_HANDLE(h);
  B2F_NEED_PATTERN(h);
  B2F_ON_DEVICE(h);
  if (dof_index < 0 || dof_index >= h->bs || n < 0 || (n > 0 && !node_ids))
    return fail(B2F_ERR_INVALID_ARGUMENT, "b2f_dirichlet_flag_penalty: bad arguments (dof_index=%d, n=%lld)", dof_index, (long long)n);
  return dirichlet_flag_penalty(h, node_ids, n, dof_index, value, penalty, 0, (cudaStream_t)stream);
}

int b2f_dirichlet_flag_weak_penalty(b2f_handle h, const int32_t* node_ids, int64_t n, int32_t dof_index, double value, double penalty, b2f_stream stream)
{
  B2F_CHECK_HANDLE(h);
  B2F_NEED_PATTERN(h);
  B2F_ON_DEVICE(h);
  if (dof_index < 0 || dof_index >= h->bs || n < 0 || (n > 0 && !node_ids))
    return fail(B2F_ERR_INVALID_ARGUMENT, "b2f_dirichlet_flag_weak_penalty: bad arguments (dof_index=%d, n=%lld)", dof_index, (long long)n);
  return dirichlet_flag_penalty(h, node_ids, n, dof_index, value, penalty, 1, (cudaStream_t)stream);
}

int b2f_dirichlet_flag_elimination(b2f_handle h, const int32_t* node_ids, int64_t n, int32_t dof_index, double value, int32_t mode, b2f_stream stream)
{
  B2F_CHECK_HANDLE(h);
  B2F_NEED_PATTERN(h);
  B2F_ON_DEVICE(h);
  if (dof_index < 0 || dof_index >= h->bs || n < 0 || (n > 0 && !node_ids) || (mode != B2F_ELIMINATE_ROW && mode != B2F_ELIMINATE_ROW_COLUMN))
    return fail(B2F_ERR_INVALID_ARGUMENT, "b2f_dirichlet_flag_elimination: bad arguments (dof_index=%d, n=%lld, mode=%d)", dof_index, (long long)n, mode);
  return dirichlet_flag_elimination(h, node_ids, n, dof_index, value, mode, (cudaStream_t)stream);
}

int b2f_dirichlet_apply(b2f_handle h, b2f_stream stream)
{
  B2F_CHECK_HANDLE(h);
  B2F_NEED_PATTERN(h);
  B2F_ON_DEVICE(h);
  return dirichlet_apply(h, (cudaStream_t)stream);
}

int b2f_spmv(b2f_handle h, const double* x, double* y, b2f_stream stream)
{
  B2F_CHECK_HANDLE(h);
  B2F_NEED_PATTERN(h);
  B2F_ON_DEVICE(h);
  if (!x || !y)
    return fail(B2F_ERR_INVALID_ARGUMENT, "b2f_spmv: null vector");
  return spmv(h, x, y, (cudaStream_t)stream);
}

int b2f_assemble_host(b2f_handle h, const b2f_operator* op, const double* coords_host, double* values_host, b2f_stream stream)
{
  B2F_CHECK_HANDLE(h);
  B2F_NEED_PATTERN(h);
  B2F_ON_DEVICE(h);
  if (!op || !coords_host)
    return fail(B2F_ERR_INVALID_ARGUMENT, "b2f_assemble_host: null argument");
  cudaStream_t s = (cudaStream_t)stream;
  const long long nn = h->mesh.nb_node;
  const size_t cbytes = sizeof(double) * 3 * (size_t)nn;
  B2F_CUDA(h->coords_stage.alloc(cbytes));
  const int variant = h->variant == B2F_ASM_DEFAULT ? B2F_ASM_NODE_GATHER : h->variant;
  const bool pipelined = variant == B2F_ASM_NODE_GATHER && h->fan_ok && nn >= (1ll << 16) && !h->reordered; // row chunks follow the caller's row order
  if (!pipelined) {
    B2F_CUDA(cudaMemcpyAsync(h->coords_stage.p, coords_host, cbytes, cudaMemcpyHostToDevice, s));
    const double* saved = h->mesh.coords;
    h->mesh.coords = h->coords_stage.as<double>();
    int rc = assemble(h, op, s);
    h->mesh.coords = saved;
    if (rc != B2F_OK)
      return rc;
    if (values_host)
      B2F_CUDA(cudaMemcpyAsync(values_host, h->values.p, sizeof(double) * (size_t)h->nb_nz, cudaMemcpyDeviceToHost, s));
    B2F_CUDA(cudaStreamSynchronize(s));
    return B2F_OK;
  }
  // Pipeline over row chunks: coordinates go up as a growing prefix on one copy stream, each chunk of rows is assembled
  // as soon as the coordinates it reads have landed, and its values go back on a second copy stream while the next
  // chunk is assembled and the next coordinates arrive (the link is full duplex).
  if (h->chunk_row.empty()) {
    int rc = pipeline_plan(h, B2F_HOST_PIPELINE_CHUNKS, s);
    if (rc != B2F_OK)
      return rc;
  }
  const int nc = (int)h->chunk_row.size() - 1;
  if (!h->pipe_h2d)
    B2F_CUDA(cudaStreamCreateWithFlags(&h->pipe_h2d, cudaStreamNonBlocking));
  if (!h->pipe_d2h)
    B2F_CUDA(cudaStreamCreateWithFlags(&h->pipe_d2h, cudaStreamNonBlocking));
  while ((int)h->pipe_events.size() < 2 * nc + 1) {
    cudaEvent_t e;
    B2F_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    h->pipe_events.push_back(e);
  }
  // work queued earlier on the caller's stream comes first
  B2F_CUDA(cudaEventRecord(h->pipe_events[2 * nc], s));
  B2F_CUDA(cudaStreamWaitEvent(h->pipe_h2d, h->pipe_events[2 * nc], 0));
  B2F_CUDA(cudaStreamWaitEvent(h->pipe_d2h, h->pipe_events[2 * nc], 0));
  const double* saved = h->mesh.coords;
  h->mesh.coords = h->coords_stage.as<double>();
  long long pos = 0;
  int rc = B2F_OK;
  cudaError_t ce = cudaSuccess;
  for (int c = 0; c < nc && rc == B2F_OK && ce == cudaSuccess; ++c) {
    const long long need = (c == nc - 1) ? nn : std::min<long long>(nn, h->chunk_node_hi[c] + 1);
    if (need > pos) {
      ce = cudaMemcpyAsync(h->coords_stage.as<double>() + 3 * pos, coords_host + 3 * pos, sizeof(double) * 3 * (size_t)(need - pos), cudaMemcpyHostToDevice, h->pipe_h2d);
      if (ce == cudaSuccess)
        ce = cudaEventRecord(h->pipe_events[2 * c], h->pipe_h2d);
      if (ce == cudaSuccess)
        ce = cudaStreamWaitEvent(s, h->pipe_events[2 * c], 0);
      pos = need;
      if (ce != cudaSuccess)
        break;
    }
    h->range_first_row = h->chunk_row[c];
    h->range_rows = h->chunk_row[c + 1] - h->chunk_row[c];
    rc = assemble(h, op, s);
    if (rc == B2F_OK && values_host) {
      const int64_t v0 = h->chunk_value_off[c], v1 = h->chunk_value_off[c + 1];
      ce = cudaEventRecord(h->pipe_events[2 * c + 1], s);
      if (ce == cudaSuccess)
        ce = cudaStreamWaitEvent(h->pipe_d2h, h->pipe_events[2 * c + 1], 0);
      if (ce == cudaSuccess && v1 > v0)
        ce = cudaMemcpyAsync(values_host + v0, h->values.as<double>() + v0, sizeof(double) * (size_t)(v1 - v0), cudaMemcpyDeviceToHost, h->pipe_d2h);
    }
  }
  h->range_rows = -1;
  h->range_first_row = 0;
  h->mesh.coords = saved;
  cudaError_t e1 = cudaStreamSynchronize(h->pipe_h2d), e2 = cudaStreamSynchronize(s), e3 = cudaStreamSynchronize(h->pipe_d2h);
  if (rc != B2F_OK)
    return rc;
  B2F_CUDA(ce);
  B2F_CUDA(e1);
  B2F_CUDA(e2);
  B2F_CUDA(e3);
  return B2F_OK;
}

int b2f_ghost_plan(b2f_handle h, void* nccl_comm, int32_t rank, int32_t world, const int32_t* node_owner_rank_host, const int64_t* node_global_id_host,
                   b2f_stream stream)
{
  B2F_CHECK_HANDLE(h);
  B2F_NEED_PATTERN(h);
  B2F_ON_DEVICE(h);
  if (h->mesh.node_is_own)
    return fail(B2F_ERR_STATE, "b2f_ghost_plan: the exchange mode assembles every local row; mesh.node_is_own must be NULL");
  return ghost_plan(h, nccl_comm, rank, world, node_owner_rank_host, node_global_id_host, (cudaStream_t)stream);
}

int b2f_ghost_exchange(b2f_handle h, void* nccl_comm, b2f_stream stream)
{
  B2F_CHECK_HANDLE(h);
  B2F_NEED_PATTERN(h);
  B2F_ON_DEVICE(h);
  return ghost_exchange(h, nccl_comm, (cudaStream_t)stream);
}

int b2f_assemble_exchange(b2f_handle h, const b2f_operator* op, void* nccl_comm, b2f_stream stream)
{
  B2F_CHECK_HANDLE(h);
  B2F_NEED_PATTERN(h);
  B2F_ON_DEVICE(h);
  if (!op)
    return fail(B2F_ERR_INVALID_ARGUMENT, "b2f_assemble_exchange: null operator");
  cudaStream_t s = (cudaStream_t)stream;
  const GhostPlan& P = h->ghost;
  if (!P.ready)
    return fail(B2F_ERR_STATE, "b2f_assemble_exchange: b2f_ghost_plan has not been called");
  const long long nn = h->mesh.nb_node;
  const int variant = h->variant == B2F_ASM_DEFAULT ? B2F_ASM_NODE_GATHER : h->variant;
  // rows the exchange reads, widened to whole tiles of the gather kernel (128 rows is a multiple of every tile height)
  const long long lo = (P.send_row_lo / 128) * 128;
  const long long hi = std::min<long long>(nn, ((P.send_row_hi + 127) / 128) * 128);
  const bool overlap = variant == B2F_ASM_NODE_GATHER && h->fan_ok && !op->accumulate && (P.total_send > 0 ? (hi - lo) * 4 <= nn : P.total_recv > 0);
  if (!overlap) { // nothing to hide the exchange behind (or the sent rows are spread over the whole numbering)
    int rc = assemble(h, op, s);
    return rc != B2F_OK ? rc : ghost_exchange(h, nccl_comm, s);
  }
  if (!h->comm_stream) {
    int least = 0, greatest = 0;
    B2F_CUDA(cudaDeviceGetStreamPriorityRange(&least, &greatest));
    B2F_CUDA(cudaStreamCreateWithPriority(&h->comm_stream, cudaStreamNonBlocking, greatest));
    for (cudaEvent_t& e : h->comm_events)
      B2F_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
  }
  // 1. the rows that travel; 2. their pack + send/recv on the communication stream; 3. every other row meanwhile;
  // 4. the received partial sums are added once both are done
  auto assemble_rows = [&](long long first, long long count, bool keep_counters) -> int {
    if (count <= 0)
      return B2F_OK;
    h->range_first_row = first;
    h->range_rows = count;
    h->range_keep_counters = keep_counters;
    const int rc = assemble(h, op, s);
    h->range_rows = -1;
    h->range_first_row = 0;
    h->range_keep_counters = false;
    return rc;
  };
  B2F_CUDA(cudaMemsetAsync(h->counters.p, 0, sizeof(unsigned long long) * 2, s));
  int rc = P.total_send > 0 ? assemble_rows(lo, hi - lo, true) : B2F_OK; // a rank that only receives posts its receives at once
  if (rc != B2F_OK)
    return rc;
  B2F_CUDA(cudaEventRecord(h->comm_events[0], s));
  B2F_CUDA(cudaStreamWaitEvent(h->comm_stream, h->comm_events[0], 0));
  rc = ghost_send_recv(h, nccl_comm, h->comm_stream);
  if (P.total_send > 0) {
    if (rc == B2F_OK)
      rc = assemble_rows(0, lo, true);
    if (rc == B2F_OK)
      rc = assemble_rows(hi, nn - hi, true);
  }
  else if (rc == B2F_OK)
    rc = assemble_rows(0, nn, true);
  // the caller's stream always rejoins the communication stream, also on an error path
  cudaError_t e1 = cudaEventRecord(h->comm_events[1], h->comm_stream);
  cudaError_t e2 = e1 == cudaSuccess ? cudaStreamWaitEvent(s, h->comm_events[1], 0) : e1;
  if (rc != B2F_OK)
    return rc;
  B2F_CUDA(e1);
  B2F_CUDA(e2);
  return ghost_add(h, s);
}

// Multi-rank twin of b2f_assemble_host: the chunk pipeline of that entry point with the ghost-row exchange of
// b2f_assemble_exchange inside it.  Chunks are assembled in row order as their coordinates land; as soon as the chunk that
// holds the last sent row is done the exchange starts on the handle's communication stream; chunks whose rows neither travel
// nor receive partial sums go back to the host at once, the others after the received sums have been added.
int b2f_assemble_exchange_host(b2f_handle h, const b2f_operator* op, void* nccl_comm, const double* coords_host, double* values_host, b2f_stream stream)
{
  B2F_CHECK_HANDLE(h);
  B2F_NEED_PATTERN(h);
  B2F_ON_DEVICE(h);
  if (!op || !coords_host)
    return fail(B2F_ERR_INVALID_ARGUMENT, "b2f_assemble_exchange_host: null argument");
  cudaStream_t s = (cudaStream_t)stream;
  const GhostPlan& P = h->ghost;
  if (!P.ready)
    return fail(B2F_ERR_STATE, "b2f_assemble_exchange_host: b2f_ghost_plan has not been called");
  const long long nn = h->mesh.nb_node;
  const size_t cbytes = sizeof(double) * 3 * (size_t)nn;
  B2F_CUDA(h->coords_stage.alloc(cbytes));
  const int variant = h->variant == B2F_ASM_DEFAULT ? B2F_ASM_NODE_GATHER : h->variant;
  const bool pipelined = variant == B2F_ASM_NODE_GATHER && h->fan_ok && nn >= (1ll << 16) && !h->reordered && !op->accumulate;
  const double* saved = h->mesh.coords;
  if (!pipelined) {
    B2F_CUDA(cudaMemcpyAsync(h->coords_stage.p, coords_host, cbytes, cudaMemcpyHostToDevice, s));
    h->mesh.coords = h->coords_stage.as<double>();
    int rc = b2f_assemble_exchange(h, op, nccl_comm, stream);
    h->mesh.coords = saved;
    if (rc != B2F_OK)
      return rc;
    if (values_host)
      B2F_CUDA(cudaMemcpyAsync(values_host, h->values.p, sizeof(double) * (size_t)h->nb_nz, cudaMemcpyDeviceToHost, s));
    B2F_CUDA(cudaStreamSynchronize(s));
    return B2F_OK;
  }
  if (h->chunk_row.empty()) {
    int rc = pipeline_plan(h, B2F_HOST_PIPELINE_CHUNKS, s);
    if (rc != B2F_OK)
      return rc;
  }
  const int nc = (int)h->chunk_row.size() - 1;
  if (!h->pipe_h2d)
    B2F_CUDA(cudaStreamCreateWithFlags(&h->pipe_h2d, cudaStreamNonBlocking));
  if (!h->pipe_d2h)
    B2F_CUDA(cudaStreamCreateWithFlags(&h->pipe_d2h, cudaStreamNonBlocking));
  while ((int)h->pipe_events.size() < 2 * nc + 1) {
    cudaEvent_t e;
    B2F_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    h->pipe_events.push_back(e);
  }
  if (!h->comm_stream) {
    int least = 0, greatest = 0;
    B2F_CUDA(cudaDeviceGetStreamPriorityRange(&least, &greatest));
    B2F_CUDA(cudaStreamCreateWithPriority(&h->comm_stream, cudaStreamNonBlocking, greatest));
    for (cudaEvent_t& e : h->comm_events)
      B2F_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
  }
  B2F_CUDA(cudaEventRecord(h->pipe_events[2 * nc], s));
  B2F_CUDA(cudaStreamWaitEvent(h->pipe_h2d, h->pipe_events[2 * nc], 0));
  B2F_CUDA(cudaStreamWaitEvent(h->pipe_d2h, h->pipe_events[2 * nc], 0));
  B2F_CUDA(cudaMemsetAsync(h->counters.p, 0, sizeof(unsigned long long) * 2, s));
  h->mesh.coords = h->coords_stage.as<double>();
  // a chunk waits for the exchange when its rows travel (they are zeroed by the pack) or receive partial sums
  auto deferred = [&](int c) {
    const bool sends = P.total_send > 0 && h->chunk_row[c] < P.send_row_hi && h->chunk_row[c + 1] > P.send_row_lo;
    const bool receives = P.total_recv > 0 && h->chunk_value_off[c] <= P.recv_val_hi && h->chunk_value_off[c + 1] > P.recv_val_lo;
    return sends || receives;
  };
  auto read_back = [&](int c) -> cudaError_t {
    const int64_t v0 = h->chunk_value_off[c], v1 = h->chunk_value_off[c + 1];
    return v1 > v0 ? cudaMemcpyAsync(values_host + v0, h->values.as<double>() + v0, sizeof(double) * (size_t)(v1 - v0), cudaMemcpyDeviceToHost, h->pipe_d2h) : cudaSuccess;
  };
  int rc = B2F_OK;
  cudaError_t ce = cudaSuccess;
  bool exchange_started = false;
  auto start_exchange = [&]() {
    ce = cudaEventRecord(h->comm_events[0], s);
    if (ce == cudaSuccess)
      ce = cudaStreamWaitEvent(h->comm_stream, h->comm_events[0], 0);
    if (ce == cudaSuccess)
      rc = ghost_send_recv(h, nccl_comm, h->comm_stream);
    exchange_started = true;
  };
  if (P.total_send == 0)
    start_exchange(); // a rank that only receives posts its receives at once
  long long pos = 0;
  for (int c = 0; c < nc && rc == B2F_OK && ce == cudaSuccess; ++c) {
    const long long need = (c == nc - 1) ? nn : std::min<long long>(nn, h->chunk_node_hi[c] + 1);
    if (need > pos) {
      ce = cudaMemcpyAsync(h->coords_stage.as<double>() + 3 * pos, coords_host + 3 * pos, sizeof(double) * 3 * (size_t)(need - pos), cudaMemcpyHostToDevice, h->pipe_h2d);
      if (ce == cudaSuccess)
        ce = cudaEventRecord(h->pipe_events[2 * c], h->pipe_h2d);
      if (ce == cudaSuccess)
        ce = cudaStreamWaitEvent(s, h->pipe_events[2 * c], 0);
      pos = need;
      if (ce != cudaSuccess)
        break;
    }
    h->range_first_row = h->chunk_row[c];
    h->range_rows = h->chunk_row[c + 1] - h->chunk_row[c];
    h->range_keep_counters = true;
    rc = assemble(h, op, s);
    if (rc != B2F_OK)
      break;
    if (!exchange_started && h->chunk_row[c + 1] >= P.send_row_hi)
      start_exchange();
    if (rc == B2F_OK && ce == cudaSuccess && values_host && !deferred(c)) {
      ce = cudaEventRecord(h->pipe_events[2 * c + 1], s);
      if (ce == cudaSuccess)
        ce = cudaStreamWaitEvent(h->pipe_d2h, h->pipe_events[2 * c + 1], 0);
      if (ce == cudaSuccess)
        ce = read_back(c);
    }
  }
  h->range_rows = -1;
  h->range_first_row = 0;
  h->range_keep_counters = false;
  h->mesh.coords = saved;
  // the caller's stream always rejoins the communication stream, also on an error path
  if (exchange_started) {
    cudaError_t e1 = cudaEventRecord(h->comm_events[1], h->comm_stream);
    if (e1 == cudaSuccess)
      e1 = cudaStreamWaitEvent(s, h->comm_events[1], 0);
    if (ce == cudaSuccess)
      ce = e1;
  }
  if (rc == B2F_OK && ce == cudaSuccess)
    rc = ghost_add(h, s);
  if (rc == B2F_OK && ce == cudaSuccess && values_host) {
    ce = cudaEventRecord(h->pipe_events[2 * nc], s);
    if (ce == cudaSuccess)
      ce = cudaStreamWaitEvent(h->pipe_d2h, h->pipe_events[2 * nc], 0);
    for (int c = 0; c < nc && ce == cudaSuccess; ++c)
      if (deferred(c))
        ce = read_back(c);
  }
  cudaError_t e1 = cudaStreamSynchronize(h->pipe_h2d), e2 = cudaStreamSynchronize(h->comm_stream), e3 = cudaStreamSynchronize(s), e4 = cudaStreamSynchronize(h->pipe_d2h);
  if (rc != B2F_OK)
    return rc;
  B2F_CUDA(ce);
  B2F_CUDA(e1);
  B2F_CUDA(e2);
  B2F_CUDA(e3);
  B2F_CUDA(e4);
  return B2F_OK;
}

} // extern "C"
