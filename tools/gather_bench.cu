// gather_bench.cu — micro-benchmark behind the k_describe design choice: the steered rBRIEF
// gather (512 byte fetches per keypoint inside a rotated 37 x 37 window) from
//   (A) the pitch-linear blurred level with LDG.U8 (what k_describe does),
//   (B) a block-linear layered CUDA array read with surf2DLayeredread (SULD, integer coordinates),
//   (C) the same array through a texture object, point sampling, unnormalised float coordinates (TEX).
// One warp per keypoint, lanes = descriptor bytes, the real rBRIEF pattern, the kernel's non-fused rotation.
// Same keypoints (random centres >= 19 px from the border, random angles) in all three; the descriptors must agree.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/build/gather_bench tools/gather_bench.cu
//   tools/build/gather_bench > gpurun_out/gather_bench.json
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { fprintf(stderr, "%s:%d %s\n", __FILE__, __LINE__, cudaGetErrorString(e_)); exit(1); } } while (0)

static __device__ __constant__ int8_t kPat[1024] = {
#include "../relative-pose-estimation_b200/csrc/orb_pattern.inc"
};

constexpr int W = 1280, H = 720, PITCH = 1280, NIMG = 64, NKP = 5000;

struct Kp { int cx, cy; float a, b; };

template <int MODE>
__global__ void __launch_bounds__(128, 8)
k_gather(const uint8_t* __restrict__ lin, cudaSurfaceObject_t surf, cudaTextureObject_t tex, const Kp* __restrict__ kps,
         uint8_t* __restrict__ desc) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int img = blockIdx.y;
  uint32_t pat[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) pat[j] = reinterpret_cast<const uint32_t*>(kPat)[lane * 8 + j];
  for (int k = blockIdx.x * 32 + warp; k < min(NKP, blockIdx.x * 32 + 32); k += 4) {
    const Kp kp = kps[img * NKP + k];
    const uint8_t* c = lin + (size_t)img * PITCH * H + (size_t)kp.cy * PITCH + kp.cx;
    uint32_t byte = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const uint32_t t = pat[j];
      const float xa = (float)(int8_t)(t & 0xff), ya = (float)(int8_t)((t >> 8) & 0xff);
      const float xb = (float)(int8_t)((t >> 16) & 0xff), yb = (float)(int8_t)(t >> 24);
      const float dxa = __fsub_rn(__fmul_rn(xa, kp.a), __fmul_rn(ya, kp.b)), dya = __fadd_rn(__fmul_rn(xa, kp.b), __fmul_rn(ya, kp.a));
      const float dxb = __fsub_rn(__fmul_rn(xb, kp.a), __fmul_rn(yb, kp.b)), dyb = __fadd_rn(__fmul_rn(xb, kp.b), __fmul_rn(yb, kp.a));
      int va, vb;
      if (MODE == 0) {
        const int ixa = __float2int_rn(dxa), iya = __float2int_rn(dya), ixb = __float2int_rn(dxb), iyb = __float2int_rn(dyb);
        va = c[iya * PITCH + ixa]; vb = c[iyb * PITCH + ixb];
      } else if (MODE == 1) {
        // rint by the 1.5 * 2^23 trick: the low mantissa bits of (d + M) are rint(d) in two's complement
        const int M = 0x4B400000;
        const int ixa = __float_as_int(__fadd_rn(dxa, 12582912.0f)) + (kp.cx - M), iya = __float_as_int(__fadd_rn(dya, 12582912.0f)) + (kp.cy - M);
        const int ixb = __float_as_int(__fadd_rn(dxb, 12582912.0f)) + (kp.cx - M), iyb = __float_as_int(__fadd_rn(dyb, 12582912.0f)) + (kp.cy - M);
        va = surf2DLayeredread<unsigned char>(surf, ixa, iya, img);
        vb = surf2DLayeredread<unsigned char>(surf, ixb, iyb, img);
      } else if (MODE == 3 || MODE == 4) {
        // strip-major layout: vertical strips SW px wide, a 128-byte line = 128 / SW rows of one strip
        constexpr int SW = MODE == 3 ? 16 : 32, LG = MODE == 3 ? 4 : 5;
        const int M = 0x4B400000;
        const int Xa = __float_as_int(__fadd_rn(dxa, 12582912.0f)) + (kp.cx - M), Ya = __float_as_int(__fadd_rn(dya, 12582912.0f)) + (kp.cy - M);
        const int Xb = __float_as_int(__fadd_rn(dxb, 12582912.0f)) + (kp.cx - M), Yb = __float_as_int(__fadd_rn(dyb, 12582912.0f)) + (kp.cy - M);
        const uint8_t* base = lin + (size_t)img * PITCH * H;
        va = base[(Xa >> LG) * (H * SW) + Ya * SW + (Xa & (SW - 1))];
        vb = base[(Xb >> LG) * (H * SW) + Yb * SW + (Xb & (SW - 1))];
      } else {
        const float fx = (float)kp.cx + 0.5f, fy = (float)kp.cy + 0.5f;
        const float rxa = __fadd_rn(__fadd_rn(dxa, 12582912.0f), -12582912.0f), rya = __fadd_rn(__fadd_rn(dya, 12582912.0f), -12582912.0f);
        const float rxb = __fadd_rn(__fadd_rn(dxb, 12582912.0f), -12582912.0f), ryb = __fadd_rn(__fadd_rn(dyb, 12582912.0f), -12582912.0f);
        va = tex2DLayered<unsigned char>(tex, rxa + fx, rya + fy, img);
        vb = tex2DLayered<unsigned char>(tex, rxb + fx, ryb + fy, img);
      }
      byte |= (uint32_t)(va < vb) << j;
    }
    desc[((size_t)img * NKP + k) * 32 + lane] = (uint8_t)byte;
  }
}

// surface write of a pitch-linear image set into the layered array, 4 px per thread (what k_blur's store would become)
__global__ void k_fill(const uint8_t* __restrict__ lin, cudaSurfaceObject_t surf) {
  const int x4 = (blockIdx.x * blockDim.x + threadIdx.x) * 4, y = blockIdx.y, img = blockIdx.z;
  if (x4 >= W) return;
  const uint32_t v = *reinterpret_cast<const uint32_t*>(lin + (size_t)img * PITCH * H + (size_t)y * PITCH + x4);
  surf2DLayeredwrite(v, surf, x4, y, img);
}
template <int SW, int LG>
__global__ void k_to_strips(const uint8_t* __restrict__ lin, uint8_t* __restrict__ out) {
  const int x4 = (blockIdx.x * blockDim.x + threadIdx.x) * 4, y = blockIdx.y, img = blockIdx.z;
  if (x4 >= W) return;
  const size_t o = (size_t)img * PITCH * H;
  *reinterpret_cast<uint32_t*>(out + o + (x4 >> LG) * (H * SW) + y * SW + (x4 & (SW - 1))) =
      *reinterpret_cast<const uint32_t*>(lin + o + (size_t)y * PITCH + x4);
}
__global__ void k_copy_lin(const uint8_t* __restrict__ lin, uint8_t* __restrict__ out) {
  const int x4 = (blockIdx.x * blockDim.x + threadIdx.x) * 4, y = blockIdx.y, img = blockIdx.z;
  if (x4 >= W) return;
  const size_t o = (size_t)img * PITCH * H + (size_t)y * PITCH + x4;
  *reinterpret_cast<uint32_t*>(out + o) = *reinterpret_cast<const uint32_t*>(lin + o);
}

int main() {
  std::vector<uint8_t> h_img((size_t)NIMG * PITCH * H);
  uint32_t s = 12345;
  auto rnd = [&]() { s = s * 1664525u + 1013904223u; return s >> 8; };
  for (auto& p : h_img) p = (uint8_t)(rnd() & 0xff);
  std::vector<Kp> h_kp((size_t)NIMG * NKP);
  for (auto& k : h_kp) {
    k.cx = 19 + (int)(rnd() % (W - 38)); k.cy = 19 + (int)(rnd() % (H - 38));
    const double th = (rnd() % 36000) * 0.01 * 3.14159265358979 / 180.0;
    k.a = (float)cos(th); k.b = (float)sin(th);
  }
  uint8_t *d_img, *d_img2, *d_desc[3];
  Kp* d_kp;
  CK(cudaMalloc(&d_img, h_img.size())); CK(cudaMalloc(&d_img2, h_img.size()));
  CK(cudaMemcpy(d_img, h_img.data(), h_img.size(), cudaMemcpyHostToDevice));
  CK(cudaMalloc(&d_kp, h_kp.size() * sizeof(Kp)));
  CK(cudaMemcpy(d_kp, h_kp.data(), h_kp.size() * sizeof(Kp), cudaMemcpyHostToDevice));
  for (auto& d : d_desc) { CK(cudaMalloc(&d, (size_t)NIMG * NKP * 32)); CK(cudaMemset(d, 0, (size_t)NIMG * NKP * 32)); }
  cudaArray_t arr;
  cudaChannelFormatDesc cd = cudaCreateChannelDesc<unsigned char>();
  CK(cudaMalloc3DArray(&arr, &cd, make_cudaExtent(W, H, NIMG), cudaArrayLayered | cudaArraySurfaceLoadStore));
  cudaResourceDesc rd = {}; rd.resType = cudaResourceTypeArray; rd.res.array.array = arr;
  cudaSurfaceObject_t surf; CK(cudaCreateSurfaceObject(&surf, &rd));
  cudaTextureDesc td = {}; td.addressMode[0] = td.addressMode[1] = cudaAddressModeClamp; td.filterMode = cudaFilterModePoint;
  td.readMode = cudaReadModeElementType; td.normalizedCoords = 0;
  cudaTextureObject_t tex; CK(cudaCreateTextureObject(&tex, &rd, &td, nullptr));
  cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  auto timed = [&](auto fn, int reps) { fn(); CK(cudaDeviceSynchronize()); CK(cudaEventRecord(e0)); for (int i = 0; i < reps; ++i) fn(); CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1)); float ms; CK(cudaEventElapsedTime(&ms, e0, e1)); CK(cudaGetLastError()); return ms / reps; };
  dim3 fg((W / 4 + 127) / 128, H, NIMG);
  const float ms_fill = timed([&] { k_fill<<<fg, 128>>>(d_img, surf); }, 10);
  const float ms_copy = timed([&] { k_copy_lin<<<fg, 128>>>(d_img, d_img2); }, 10);
  dim3 grid((NKP + 31) / 32, NIMG);
  const float ms0 = timed([&] { k_gather<0><<<grid, 128>>>(d_img, surf, tex, d_kp, d_desc[0]); }, 10);
  const float ms1 = timed([&] { k_gather<1><<<grid, 128>>>(d_img, surf, tex, d_kp, d_desc[1]); }, 10);
  const float ms2 = timed([&] { k_gather<2><<<grid, 128>>>(d_img, surf, tex, d_kp, d_desc[2]); }, 10);
  uint8_t *d_s16, *d_s32, *d_desc34[2];
  CK(cudaMalloc(&d_s16, h_img.size())); CK(cudaMalloc(&d_s32, h_img.size()));
  for (auto& d : d_desc34) { CK(cudaMalloc(&d, (size_t)NIMG * NKP * 32)); CK(cudaMemset(d, 0, (size_t)NIMG * NKP * 32)); }
  const float ms_s16 = timed([&] { k_to_strips<16, 4><<<fg, 128>>>(d_img, d_s16); }, 10);
  const float ms_s32 = timed([&] { k_to_strips<32, 5><<<fg, 128>>>(d_img, d_s32); }, 10);
  const float ms3 = timed([&] { k_gather<3><<<grid, 128>>>(d_s16, surf, tex, d_kp, d_desc34[0]); }, 10);
  const float ms4 = timed([&] { k_gather<4><<<grid, 128>>>(d_s32, surf, tex, d_kp, d_desc34[1]); }, 10);
  std::vector<uint8_t> h3((size_t)NIMG * NKP * 32), h4(h3.size());
  CK(cudaMemcpy(h3.data(), d_desc34[0], h3.size(), cudaMemcpyDeviceToHost));
  CK(cudaMemcpy(h4.data(), d_desc34[1], h3.size(), cudaMemcpyDeviceToHost));
  std::vector<uint8_t> h0((size_t)NIMG * NKP * 32), h1(h0.size()), h2(h0.size());
  CK(cudaMemcpy(h0.data(), d_desc[0], h0.size(), cudaMemcpyDeviceToHost));
  CK(cudaMemcpy(h1.data(), d_desc[1], h0.size(), cudaMemcpyDeviceToHost));
  CK(cudaMemcpy(h2.data(), d_desc[2], h0.size(), cudaMemcpyDeviceToHost));
  size_t d01 = 0, d02 = 0;
  for (size_t i = 0; i < h0.size(); ++i) { d01 += h0[i] != h1[i]; d02 += h0[i] != h2[i]; }
  size_t d03 = 0, d04 = 0;
  for (size_t i = 0; i < h0.size(); ++i) { d03 += h0[i] != h3[i]; d04 += h0[i] != h4[i]; }
  printf("{\"ms_ldg_strips16x8\": %.4f, \"ms_ldg_strips32x4\": %.4f, \"ms_store_strips16\": %.4f, \"ms_store_strips32\": %.4f, "
         "\"descriptor_bytes_differing_strips\": [%zu, %zu]}\n", ms3, ms4, ms_s16, ms_s32, d03, d04);
  printf("{\"images\": %d, \"keypoints_per_image\": %d, \"ms_ldg_pitch_linear\": %.4f, \"ms_suld_layered_array\": %.4f, "
         "\"ms_tex_layered_array\": %.4f, \"ms_surface_fill_4px_per_thread\": %.4f, \"ms_linear_copy_4px_per_thread\": %.4f, "
         "\"descriptor_bytes_differing_suld\": %zu, \"descriptor_bytes_differing_tex\": %zu}\n",
         NIMG, NKP, ms0, ms1, ms2, ms_fill, ms_copy, d01, d02);
  return 0;
}
