// boundary.cu — RHS and Dirichlet kernels on the device-resident system (sm_100a).
//
//   rhs_constant_source         BoundaryCondtionsHelpers::applyConstantSourceToRhsBase, femutils/ArcaneFemFunctionsGpu.h:402-429
//   rhs_nodal_source            transient heat source (T_old / dt) * |cell| / NV, modules/heat/FemModule.cc:480-495
//   rhs_neumann                 BoundaryConditions{2D,3D}::applyNeumannToRhs, :621-677 / :711-768 (geometry :88-197)
//   add_face_mass               convection boundary term of the heat module, modules/heat/FemModule.cc:628-672,813-840
//   dirichlet_flag_penalty      applyDirichletToNodeGroupViaPenalty, :434-456
//   dirichlet_flag_elimination  applyDirichletToNodeGroupViaRowElimination<MODE>, :461-482
//   dirichlet_apply             HypreDoFLinearSystemImpl::_applyRowElimination / _applyForcedValuesToLhs,
//                               femutils/HypreDoFLinearSystem.cc:319-382, and the row+column elimination semantics of
//                               femutils/AlephDoFLinearSystem.cc:501-583 (CPU-only in the reference).
// The matrix kernels work on the BSR storage directly (both value layouts), so no CSR copy is needed first.
#include "internal.h"
#include "b2f/assemble.cuh"

namespace b2f {

namespace {

  inline unsigned grid_for(long long n, int threads) { return (unsigned)((n + threads - 1) / threads); }

  __global__ void k_fill_f64(double* p, long long n, double v)
  {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n)
      p[i] = v;
  }

  template <int NV>
  __global__ void __launch_bounds__(256) k_rhs_source(const int32_t* __restrict__ cell_node, const double* __restrict__ coords, const uint8_t* __restrict__ own,
                                                      long long nb_cell, int bs, double qdot, double* __restrict__ rhs)
  {
    const long long cell = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (cell >= nb_cell)
      return;
    int32_t n[NV];
    load_cell_nodes<NV>(cell_node, cell, n);
    double X[NV][3];
#pragma unroll
    for (int k = 0; k < NV; ++k)
      load_xyz(coords, n[k], X[k]);
    const double ax = X[1][0] - X[0][0], ay = X[1][1] - X[0][1], az = X[1][2] - X[0][2];
    const double bx = X[2][0] - X[0][0], by = X[2][1] - X[0][1], bz = X[2][2] - X[0][2];
    const double cx = ay * bz - az * by, cy = az * bx - ax * bz, cz = ax * by - ay * bx;
    double domain;
    if (NV == 3)
      domain = sqrt(cx * cx + cy * cy + cz * cz) / 2.0; // computeAreaTria3, :75-83
    else {
      // computeVolumeTetra4, :110-122: |v0 . (v1 x v2)| / 6
      const double dx = X[NV - 1][0] - X[0][0], dy = X[NV - 1][1] - X[0][1], dz = X[NV - 1][2] - X[0][2];
      const double c2x = by * dz - bz * dy, c2y = bz * dx - bx * dz, c2z = bx * dy - by * dx;
      domain = fabs(ax * c2x + ay * c2y + az * c2z) / 6.0;
    }
    const double value = qdot * domain / NV;
#pragma unroll
    for (int k = 0; k < NV; ++k)
      if (own == nullptr || own[n[k]])
        atomicAdd(rhs + (long long)n[k] * bs, value);
  }

  // Nodal-field source of the transient heat module, modules/heat/FemModule.cc:480-495: rhs[n] += (field[n] * scale) * |cell| / NV
  // for own nodes that are not skipped (m_node_is_temperature_fixed).  Triangles use the module's signed 2D area
  // (_computeAreaTriangle3, :605-612); tets (3D twin) |det| / 6.
  template <int NV>
  __global__ void __launch_bounds__(256) k_rhs_nodal_source(const int32_t* __restrict__ cell_node, const double* __restrict__ coords, const uint8_t* __restrict__ own,
                                                            const uint8_t* __restrict__ skip, const double* __restrict__ field,
                                                            long long nb_cell, int bs, double scale, double* __restrict__ rhs)
  {
    const long long cell = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (cell >= nb_cell) // every cell, ghost layer included, like the reference's allCells() loop; only own nodes receive
      return;
    int32_t n[NV];
    load_cell_nodes<NV>(cell_node, cell, n);
    double X[NV][3];
#pragma unroll
    for (int k = 0; k < NV; ++k)
      load_xyz(coords, n[k], X[k]);
    double domain;
    if (NV == 3)
      domain = 0.5 * ((X[1][0] - X[0][0]) * (X[2][1] - X[0][1]) - (X[2][0] - X[0][0]) * (X[1][1] - X[0][1]));
    else {
      const double ax = X[1][0] - X[0][0], ay = X[1][1] - X[0][1], az = X[1][2] - X[0][2];
      const double bx = X[2][0] - X[0][0], by = X[2][1] - X[0][1], bz = X[2][2] - X[0][2];
      const double dx = X[NV - 1][0] - X[0][0], dy = X[NV - 1][1] - X[0][1], dz = X[NV - 1][2] - X[0][2];
      const double c2x = by * dz - bz * dy, c2y = bz * dx - bx * dz, c2z = bx * dy - by * dx;
      domain = fabs(ax * c2x + ay * c2y + az * c2z) / 6.0;
    }
#pragma unroll
    for (int k = 0; k < NV; ++k)
      if ((own == nullptr || own[n[k]]) && (skip == nullptr || !skip[n[k]]))
        atomicAdd(rhs + (long long)n[k] * bs, (field[n[k]] * scale) * domain / NV);
  }

  template <int NVF>
  __global__ void __launch_bounds__(256) k_rhs_neumann(b2f_neumann bc, const double* __restrict__ coords, const uint8_t* __restrict__ own, int bs, double* __restrict__ rhs)
  {
    const long long f = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (f >= bc.nb_face)
      return;
    int32_t n[NVF];
    double X[NVF][3];
#pragma unroll
    for (int k = 0; k < NVF; ++k) {
      n[k] = bc.face_node[f * NVF + k];
      load_xyz(coords, n[k], X[k]);
    }
    const bool outside = bc.face_outside == nullptr || bc.face_outside[f] != 0;
    double contrib;
    if (NVF == 2) {
      // computeLengthFace :132-138, computeNormalFace :149-165
      const double ex = X[1][0] - X[0][0], ey = X[1][1] - X[0][1];
      const double length = sqrt(ex * ex + ey * ey);
      if (bc.has_value)
        contrib = bc.value * length / 2.0;
      else {
        const double sx = outside ? ex : -ex, sy = outside ? ey : -ey; // (m1 - m0) after the optional swap
        const double norm = sqrt(sy * sy + sx * sx);
        const double nx = sy / norm, ny = -sx / norm;
        contrib = (nx * bc.value_x + ny * bc.value_y) * length / 2.0;
      }
    }
    else {
      // computeAreaTria :88-96, computeNormalTriangle :176-197
      const double ax = X[1][0] - X[0][0], ay = X[1][1] - X[0][1], az = X[1][2] - X[0][2];
      const double bx = X[NVF - 1][0] - X[0][0], by = X[NVF - 1][1] - X[0][1], bz = X[NVF - 1][2] - X[0][2];
      const double cx = ay * bz - az * by, cy = az * bx - ax * bz, cz = ax * by - ay * bx;
      const double area = sqrt(cx * cx + cy * cy + cz * cz) / 2.0;
      if (bc.has_value)
        contrib = bc.value * area / 3.0;
      else {
        // swapping vertex0 and vertex1: edge1 = v0 - v1, edge2 = v2 - v1
        double e1x = ax, e1y = ay, e1z = az, e2x = bx, e2y = by, e2z = bz;
        if (!outside) {
          e1x = -ax, e1y = -ay, e1z = -az;
          e2x = X[NVF - 1][0] - X[1][0], e2y = X[NVF - 1][1] - X[1][1], e2z = X[NVF - 1][2] - X[1][2];
        }
        const double nx = e1y * e2z - e1z * e2y, ny = e1z * e2x - e1x * e2z, nz = e1x * e2y - e1y * e2x;
        const double norm = sqrt(nx * nx + ny * ny + nz * nz);
        contrib = ((nx / norm) * bc.value_x + (ny / norm) * bc.value_y + (nz / norm) * bc.value_z) * area / 3.0;
      }
    }
#pragma unroll
    for (int k = 0; k < NVF; ++k)
      if (own == nullptr || own[n[k]])
        atomicAdd(rhs + (long long)n[k] * bs, contrib);
  }

  // Convection (Robin) term of the heat module, modules/heat/FemModule.cc:628-672 + :813-840: per boundary face
  // K[a,b] += h * |face| / (NVF (NVF + 1)) * (1 + delta_ab) over the face's nodes (edges: h L/6 [[2,1],[1,2]], L from x and y
  // only like _computeEdgeLength2 :617-623; triangles, the 3D twin: h A/12 (1 + delta)).  Own rows only.  Surface-sized: one
  // thread per face, NVF^2 fp64 REDs, column slot by search in the sorted row.  BS = 1 (both value layouts coincide).
  template <int NVF>
  __global__ void __launch_bounds__(256) k_add_face_mass(long long nb_face, const int32_t* __restrict__ face_node, double h, const double* __restrict__ coords,
                                                         const uint8_t* __restrict__ own, const int32_t* __restrict__ row_index, const int32_t* __restrict__ columns,
                                                         double* __restrict__ values, unsigned long long* __restrict__ counters, int diag_last)
  {
    const long long f = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (f >= nb_face)
      return;
    int32_t n[NVF];
    double X[NVF][3];
#pragma unroll
    for (int k = 0; k < NVF; ++k) {
      n[k] = face_node[f * NVF + k];
      load_xyz(coords, n[k], X[k]);
    }
    double off;
    if (NVF == 2) {
      const double ex = X[1][0] - X[0][0], ey = X[1][1] - X[0][1];
      off = (sqrt(ex * ex + ey * ey) / 3.0) * 0.5 * h;
    }
    else {
      const double ax = X[1][0] - X[0][0], ay = X[1][1] - X[0][1], az = X[1][2] - X[0][2];
      const double bx = X[NVF - 1][0] - X[0][0], by = X[NVF - 1][1] - X[0][1], bz = X[NVF - 1][2] - X[0][2];
      const double cx = ay * bz - az * by, cy = az * bx - ax * bz, cz = ax * by - ay * bx;
      off = ((sqrt(cx * cx + cy * cy + cz * cz) / 2.0) / 12.0) * h;
    }
#pragma unroll
    for (int a = 0; a < NVF; ++a) {
      if (own != nullptr && !own[n[a]])
        continue;
      const int32_t rb = __ldg(row_index + n[a]);
      const int k = __ldg(row_index + n[a] + 1) - rb;
#pragma unroll
      for (int b = 0; b < NVF; ++b) {
        const int x = (a == b) ? diag_slot(k, diag_last) : find_slot(columns + rb, k, n[b], diag_last);
        if (x < 0)
          atomicAdd(counters, 1ull);
        else
          atomicAdd(values + rb + x, (a == b) ? off * 2.0 : off);
      }
    }
  }

  __global__ void k_flag_penalty(const int32_t* __restrict__ ids, long long n, const uint8_t* __restrict__ own, int bs, int dof_index, double value, double penalty, int weak,
                                 uint8_t* __restrict__ forced_info, double* __restrict__ forced_value, double* __restrict__ rhs)
  {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n)
      return;
    const int32_t node = ids[i];
    if (own != nullptr && !own[node])
      return;
    const long long dof = (long long)node * bs + dof_index;
    forced_info[dof] = weak ? 2 : 1; // 2: WeakPenalty, a_ii += P (modules/testlab/FemModule.cc:687-717)
    forced_value[dof] = penalty;
    rhs[dof] = penalty * value;
  }

  __global__ void k_flag_elimination(const int32_t* __restrict__ ids, long long n, const uint8_t* __restrict__ own, int bs, int dof_index, double value, int mode,
                                     uint8_t* __restrict__ info, double* __restrict__ val)
  {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n)
      return;
    const int32_t node = ids[i];
    if (own != nullptr && !own[node])
      return;
    const long long dof = (long long)node * bs + dof_index;
    info[dof] = (uint8_t)mode;
    val[dof] = value;
  }

  struct SysPtrs
  {
    const int32_t* row_index;
    const int32_t* columns;
    double* values;
    const uint8_t* elim_info;
    const double* elim_value;
    const uint8_t* forced_info;
    const double* forced_value;
    double* rhs;
    long long nb_dof;
    int diag_last;
  };

  // Row+column elimination, phase 1 (AlephDoFLinearSystem.cc:540-563): rhs[c] -= A[r,c] * g_r for every stored (r,c),
  // r != c, r flagged ROW_COLUMN.  One thread per column dof c; A[r,c] is read from row r as Aleph does.
  template <int BS, bool PER_BLOCK>
  __global__ void __launch_bounds__(256) k_rowcol_rhs(SysPtrs S)
  {
    const long long c = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= S.nb_dof)
      return;
    const int32_t bc = (int32_t)(c / BS);
    const int jc = (int)(c % BS);
    const int32_t cb = S.row_index[bc], ce = S.row_index[bc + 1];
    double acc = S.rhs[c];
    bool touched = false;
    for (int32_t q = cb; q < ce; ++q) {
      const int32_t br = S.columns[q]; // pattern is symmetric: rows r that hold column c are the columns of row c
      for (int ir = 0; ir < BS; ++ir) {
        const long long r = (long long)br * BS + ir;
        if (r == c || S.elim_info[r] != B2F_ELIMINATE_ROW_COLUMN)
          continue;
        const int32_t rb = S.row_index[br];
        const int k = S.row_index[br + 1] - rb;
        const int x = (br == bc) ? diag_slot(k, S.diag_last) : find_slot(S.columns + rb, k, bc, S.diag_last);
        if (x < 0)
          continue;
        acc = acc - S.values[value_index<BS, PER_BLOCK>(rb, k, x, ir, jc)] * S.elim_value[r];
        touched = true;
      }
    }
    if (touched)
      S.rhs[c] = acc;
  }

  // One thread per row dof r, all writes stay inside row r:
  //   (a) row+column: A[r,c] = delta_rc where r or c is flagged ROW_COLUMN; rhs[r] = g_r   (AlephDoFLinearSystem.cc:565-583)
  //   (b) row elimination: row r <- unit row, rhs[r] = g_r                                   (HypreDoFLinearSystem.cc:319-351)
  //   (c) forced value: A[r,r] = P                                                            (HypreDoFLinearSystem.cc:356-382)
  template <int BS, bool PER_BLOCK>
  __global__ void __launch_bounds__(256) k_dirichlet_rows(SysPtrs S, int any_rowcol)
  {
    const long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= S.nb_dof)
      return;
    const uint8_t info = S.elim_info[r];
    const bool forced = S.forced_info[r] != 0;
    if (!any_rowcol && info == B2F_ELIMINATE_NONE && !forced)
      return;
    const int32_t br = (int32_t)(r / BS);
    const int ir = (int)(r % BS);
    const int32_t rb = S.row_index[br];
    const int k = S.row_index[br + 1] - rb;
    if (any_rowcol) {
      for (int x = 0; x < k; ++x) {
        const int32_t bcol = S.columns[rb + x];
        for (int j = 0; j < BS; ++j) {
          const long long c = (long long)bcol * BS + j;
          if (info == B2F_ELIMINATE_ROW_COLUMN || S.elim_info[c] == B2F_ELIMINATE_ROW_COLUMN)
            S.values[value_index<BS, PER_BLOCK>(rb, k, x, ir, j)] = (c == r) ? 1.0 : 0.0;
        }
      }
      if (info == B2F_ELIMINATE_ROW_COLUMN)
        S.rhs[r] = S.elim_value[r];
    }
    if (info == B2F_ELIMINATE_ROW) {
      for (int x = 0; x < k; ++x) {
        const int32_t bcol = S.columns[rb + x];
        for (int j = 0; j < BS; ++j)
          S.values[value_index<BS, PER_BLOCK>(rb, k, x, ir, j)] = ((long long)bcol * BS + j == r) ? 1.0 : 0.0;
      }
      S.rhs[r] = S.elim_value[r];
    }
    if (forced) { // the diagonal block: slot 0, or k - 1 in a diagonal-last row
      double* d = S.values + value_index<BS, PER_BLOCK>(rb, k, diag_slot(k, S.diag_last), ir, ir);
      *d = (S.forced_info[r] == 2) ? *d + S.forced_value[r] : S.forced_value[r];
    }
  }

  template <int BS, bool PER_BLOCK>
  int run_dirichlet(b2f_context_s* h, const SysPtrs& S, bool any_rowcol, cudaStream_t s)
  {
    if (any_rowcol) {
      k_rowcol_rhs<BS, PER_BLOCK><<<grid_for(S.nb_dof, 256), 256, 0, s>>>(S);
      B2F_CUDA(cudaGetLastError());
    }
    k_dirichlet_rows<BS, PER_BLOCK><<<grid_for(S.nb_dof, 256), 256, 0, s>>>(S, any_rowcol ? 1 : 0);
    B2F_CUDA(cudaGetLastError());
    (void)h;
    return B2F_OK;
  }

  __global__ void k_any_rowcol(const uint8_t* __restrict__ info, long long n, unsigned long long* __restrict__ flag)
  {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const bool hit = i < n && info[i] == B2F_ELIMINATE_ROW_COLUMN;
    if (__any_sync(0xffffffffu, hit) && (threadIdx.x & 31) == 0)
      atomicOr(flag, 1ull);
  }

} // namespace

int rhs_fill(b2f_context_s* h, double v, cudaStream_t s)
{
  if (h->nb_dof > 0)
    k_fill_f64<<<grid_for(h->nb_dof, 256), 256, 0, s>>>(h->rhs.as<double>(), h->nb_dof, v);
  B2F_CUDA(cudaGetLastError());
  return B2F_OK;
}

int clear_values(b2f_context_s* h, cudaStream_t s)
{
  // HypreDoFLinearSystemImpl::clearValues (HypreDoFLinearSystem.cc:184-191) + the callers' rhs.fill(0)
  // (modules/poisson/FemModule.cc:152-160).
  B2F_CUDA(cudaMemsetAsync(h->forced_info.p, 0, (size_t)h->nb_dof, s));
  B2F_CUDA(cudaMemsetAsync(h->elim_info.p, 0, (size_t)h->nb_dof, s));
  B2F_CUDA(cudaMemsetAsync(h->elim_value.p, 0, sizeof(double) * (size_t)h->nb_dof, s));
  B2F_CUDA(cudaMemsetAsync(h->rhs.p, 0, sizeof(double) * (size_t)h->nb_dof, s));
  return B2F_OK;
}

int rhs_constant_source(b2f_context_s* h, double qdot, cudaStream_t s)
{
  const long long nc = h->mesh.nb_cell;
  if (nc == 0)
    return B2F_OK;
  if (h->nv == 3)
    k_rhs_source<3><<<grid_for(nc, 256), 256, 0, s>>>(h->mesh.cell_node, h->mesh.coords, h->mesh.node_is_own, nc, h->bs, qdot, h->rhs.as<double>());
  else
    k_rhs_source<4><<<grid_for(nc, 256), 256, 0, s>>>(h->mesh.cell_node, h->mesh.coords, h->mesh.node_is_own, nc, h->bs, qdot, h->rhs.as<double>());
  B2F_CUDA(cudaGetLastError());
  return B2F_OK;
}

int rhs_nodal_source(b2f_context_s* h, const double* node_field, double scale, const uint8_t* node_skip, cudaStream_t s)
{
  const long long nc = h->mesh.nb_cell;
  if (nc == 0)
    return B2F_OK;
  if (h->nv == 3)
    k_rhs_nodal_source<3><<<grid_for(nc, 256), 256, 0, s>>>(h->mesh.cell_node, h->mesh.coords, h->mesh.node_is_own, node_skip, node_field, nc, h->bs,
                                                           scale, h->rhs.as<double>());
  else
    k_rhs_nodal_source<4><<<grid_for(nc, 256), 256, 0, s>>>(h->mesh.cell_node, h->mesh.coords, h->mesh.node_is_own, node_skip, node_field, nc, h->bs,
                                                           scale, h->rhs.as<double>());
  B2F_CUDA(cudaGetLastError());
  return B2F_OK;
}

int rhs_neumann(b2f_context_s* h, const b2f_neumann* bc, cudaStream_t s)
{
  if (bc->nb_face == 0)
    return B2F_OK;
  if (h->mesh.dim == 2)
    k_rhs_neumann<2><<<grid_for(bc->nb_face, 256), 256, 0, s>>>(*bc, h->mesh.coords, h->mesh.node_is_own, h->bs, h->rhs.as<double>());
  else
    k_rhs_neumann<3><<<grid_for(bc->nb_face, 256), 256, 0, s>>>(*bc, h->mesh.coords, h->mesh.node_is_own, h->bs, h->rhs.as<double>());
  B2F_CUDA(cudaGetLastError());
  return B2F_OK;
}

int add_face_mass(b2f_context_s* h, int64_t nb_face, const int32_t* face_node, double coefficient, cudaStream_t s)
{
  B2F_CUDA(cudaMemsetAsync(h->counters.p, 0, sizeof(unsigned long long), s));
  if (nb_face == 0)
    return B2F_OK;
  if (h->mesh.dim == 2)
    k_add_face_mass<2><<<grid_for(nb_face, 256), 256, 0, s>>>(nb_face, face_node, coefficient, h->mesh.coords, h->mesh.node_is_own, h->row_index.as<int32_t>(),
                                                             h->columns.as<int32_t>(), h->values.as<double>(), h->counters.as<unsigned long long>(), h->diag_last);
  else
    k_add_face_mass<3><<<grid_for(nb_face, 256), 256, 0, s>>>(nb_face, face_node, coefficient, h->mesh.coords, h->mesh.node_is_own, h->row_index.as<int32_t>(),
                                                             h->columns.as<int32_t>(), h->values.as<double>(), h->counters.as<unsigned long long>(), h->diag_last);
  B2F_CUDA(cudaGetLastError());
  return B2F_OK;
}

int dirichlet_flag_penalty(b2f_context_s* h, const int32_t* ids, int64_t n, int dof_index, double value, double penalty, int weak, cudaStream_t s)
{
  if (n == 0)
    return B2F_OK;
  k_flag_penalty<<<grid_for(n, 256), 256, 0, s>>>(ids, n, h->mesh.node_is_own, h->bs, dof_index, value, penalty, weak, h->forced_info.as<uint8_t>(), h->forced_value.as<double>(),
                                                  h->rhs.as<double>());
  B2F_CUDA(cudaGetLastError());
  return B2F_OK;
}

int dirichlet_flag_elimination(b2f_context_s* h, const int32_t* ids, int64_t n, int dof_index, double value, int mode, cudaStream_t s)
{
  if (n == 0)
    return B2F_OK;
  k_flag_elimination<<<grid_for(n, 256), 256, 0, s>>>(ids, n, h->mesh.node_is_own, h->bs, dof_index, value, mode, h->elim_info.as<uint8_t>(), h->elim_value.as<double>());
  B2F_CUDA(cudaGetLastError());
  return B2F_OK;
}

int dirichlet_apply(b2f_context_s* h, cudaStream_t s)
{
  SysPtrs S;
  S.row_index = h->row_index.as<int32_t>();
  S.columns = h->columns.as<int32_t>();
  S.values = h->values.as<double>();
  S.elim_info = h->elim_info.as<uint8_t>();
  S.elim_value = h->elim_value.as<double>();
  S.forced_info = h->forced_info.as<uint8_t>();
  S.forced_value = h->forced_value.as<double>();
  S.rhs = h->rhs.as<double>();
  S.nb_dof = h->nb_dof;
  S.diag_last = h->diag_last;
  if (S.nb_dof == 0)
    return B2F_OK;
  // Does any dof ask for row+column elimination?  (The flag arrays are caller-writable, so ask the device.)
  unsigned long long* flag = h->counters.as<unsigned long long>() + 3;
  B2F_CUDA(cudaMemsetAsync(flag, 0, sizeof(unsigned long long), s));
  k_any_rowcol<<<grid_for(S.nb_dof, 256), 256, 0, s>>>(S.elim_info, S.nb_dof, flag);
  unsigned long long host_flag = 0;
  B2F_CUDA(cudaMemcpyAsync(&host_flag, flag, sizeof(host_flag), cudaMemcpyDeviceToHost, s));
  B2F_CUDA(cudaStreamSynchronize(s));
  const bool any_rc = host_flag != 0;
  const bool pb = h->per_block;
  switch (h->bs) {
  case 1:
    return run_dirichlet<1, true>(h, S, any_rc, s);
  case 2:
    return pb ? run_dirichlet<2, true>(h, S, any_rc, s) : run_dirichlet<2, false>(h, S, any_rc, s);
  case 3:
    return pb ? run_dirichlet<3, true>(h, S, any_rc, s) : run_dirichlet<3, false>(h, S, any_rc, s);
  }
  return fail(B2F_ERR_INVALID_ARGUMENT, "unsupported block size %d", h->bs);
}

} // namespace b2f
