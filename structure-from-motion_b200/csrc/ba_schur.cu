// Camera-side reduction of the normal equations (K2b).
//
// Reference: ba_processor.py:376-382 forms A = J_p^T J_p + lambda I, e_p, B = J_p^T J_x densely and
// inverts A - B D^-1 B^T.  Here each camera-major chunk (<= 256 observations of ONE camera) reduces,
// in f64, the 63 numbers the reduced system needs from that camera:
//   A_i   = sum J_p^T J_p                      (28, upper triangle)
//   Ssub  = sum J_p^T Y J_p,  Y = J_x D^-1 J_x^T (28)   -> S_ii = A_i + lambda I - Ssub (block-Jacobi)
//   g_i   = sum J_p^T rho,    rho = r - J_x D^-1 e_x (7) -> right-hand side e_p - B D^-1 e_x
// The chunk's records are staged in shared memory as doubles; thread (entry, part) then walks 64
// observations with 4 LDS.64 + 2 DFMA each -- a skinny Gram product, no atomics, fixed order.
#include "ba_internal.h"
#include "sfmba_device.cuh"

namespace sfmba {

constexpr int ROW = 30;   // doubles per staged observation: Jp(14) T(14) rho(2)

struct Cols {
  unsigned char l0, r0, l1, r1;
};

__device__ __forceinline__ Cols entry_cols(int e) {
  // e < 28: A(a,b); e < 56: Ssub(a,b); e < 63: g(a)
  Cols c;
  int k = e < 28 ? e : e - 28;
  if (e < 56) {
    int a = 0;
    while (k >= 7 - a) {
      k -= 7 - a;
      ++a;
    }
    const int b = a + k;
    c.l0 = a;
    c.l1 = 7 + a;
    if (e < 28) {
      c.r0 = b;
      c.r1 = 7 + b;
    } else {
      c.r0 = 14 + b;
      c.r1 = 21 + b;
    }
  } else {
    const int a = e - 56;
    c.l0 = a;
    c.l1 = 7 + a;
    c.r0 = 28;
    c.r1 = 29;
  }
  return c;
}

__global__ void __launch_bounds__(256) k_camreduce(BaDev d) {
  extern __shared__ __align__(16) double s_rows[];   // [CHUNK_OBS][ROW] + [4][64]
  double* s_part = s_rows + CHUNK_OBS * ROW;
  const int c = blockIdx.x;
  if (c >= d.sc->num_chunks) return;
  const int cam = d.chunk_cam[c];
  const int beg = d.chunk_beg[c];
  int end = beg + CHUNK_OBS;
  const int cam_end = d.cam_rowptr[cam + 1];
  if (end > cam_end) end = cam_end;
  const int len = end - beg;
  const int tid = threadIdx.x;
  {
    double* row = s_rows + tid * ROW;
    if (tid < len) {
      const float4* rec = d.rec_cm + 4 * (size_t)(beg + tid);
      const float4 a = rec[0], b = rec[1], cc = rec[2], dd = rec[3];
      const float4 y = d.y_cm[beg + tid];
      const double jp[14] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w, cc.x, cc.y, cc.z, cc.w, dd.x, dd.y};
      const double yuu = y.x, yuv = y.y, yvv = y.z;
#pragma unroll
      for (int k = 0; k < 7; ++k) {
        row[k] = jp[k];
        row[7 + k] = jp[7 + k];
        row[14 + k] = yuu * jp[k] + yuv * jp[7 + k];
        row[21 + k] = yuv * jp[k] + yvv * jp[7 + k];
      }
      row[28] = dd.z;
      row[29] = dd.w;
    } else {
#pragma unroll
      for (int k = 0; k < ROW; ++k) row[k] = 0.0;
    }
  }
  __syncthreads();
  const int part = tid >> 6, e = tid & 63;
  double acc = 0.0;
  if (e < 63) {
    const Cols cl = entry_cols(e);
    const double* base = s_rows + (part * 64) * ROW;
#pragma unroll 4
    for (int o = 0; o < 64; ++o) {
      const double* row = base + o * ROW;
      acc += row[cl.l0] * row[cl.r0] + row[cl.l1] * row[cl.r1];
    }
  }
  s_part[part * 64 + e] = acc;
  __syncthreads();
  if (tid < 64) {
    const double v = (s_part[tid] + s_part[64 + tid]) + (s_part[128 + tid] + s_part[192 + tid]);
    d.chunk_acc[(size_t)c * CAM_ACC + tid] = v;
  }
}

// cam_acc[cam][e] = sum over the camera's chunks, in chunk order
__global__ void k_cam_finalize(BaDev d) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= d.V * CAM_ACC) return;
  const int cam = idx / CAM_ACC, e = idx % CAM_ACC;
  double a = 0.0;
  const int c0 = d.cam_chunk_ptr[cam], c1 = d.cam_chunk_ptr[cam + 1];
  for (int c = c0; c < c1; ++c) a += d.chunk_acc[(size_t)c * CAM_ACC + e];
  d.cam_acc[idx] = a;
}

// per camera: A_i + lambda I (full 7x7), block-Jacobi inverse of S_ii, right-hand side g_i
__global__ void k_cam_blocks(BaDev d) {
  const int cam = blockIdx.x * blockDim.x + threadIdx.x;
  if (cam >= d.V) return;
  const double lam = d.sc->lambda;
  const double* acc = d.cam_acc + (size_t)cam * CAM_ACC;
  double A[49], S[49];
  int k = 0;
  for (int a = 0; a < 7; ++a) {
    for (int b = a; b < 7; ++b, ++k) {
      double va = acc[k], vs = acc[k] - acc[28 + k];
      if (a == b) {
        va += lam;
        vs += lam;
      }
      A[a * 7 + b] = va;
      A[b * 7 + a] = va;
      S[a * 7 + b] = vs;
      S[b * 7 + a] = vs;
    }
  }
  if (!spd_inverse<7>(S)) atomicOr(&d.sc->flags, FLAG_SINGULAR);
  double* ao = d.ablk + (size_t)cam * 49;
  double* mo = d.minv + (size_t)cam * 49;
  for (int i = 0; i < 49; ++i) {
    ao[i] = A[i];
    mo[i] = S[i];
  }
  for (int a = 0; a < 7; ++a) d.gvec[7 * cam + a] = acc[56 + a];
}

void launch_camreduce(const BaDev& d, cudaStream_t st, LaunchCounter& lc) {
  const size_t smem = (size_t)(CHUNK_OBS * ROW + 4 * 64) * sizeof(double);
  static bool attr_set = false;
  if (!attr_set) {
    cudaFuncSetAttribute(k_camreduce, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    attr_set = true;
  }
  if (d.nchunk_bound > 0) {
    k_camreduce<<<d.nchunk_bound, 256, smem, st>>>(d);
    lc.launches++;
  }
}

void launch_cam_finalize(const BaDev& d, cudaStream_t st, LaunchCounter& lc) {
  const int n = d.V * CAM_ACC;
  k_cam_finalize<<<(n + 255) / 256, 256, 0, st>>>(d);
  lc.launches++;
}

void launch_cam_blocks(const BaDev& d, cudaStream_t st, LaunchCounter& lc) {
  k_cam_blocks<<<(d.V + 63) / 64, 64, 0, st>>>(d);
  lc.launches++;
}

}  // namespace sfmba
