// How fast can ONE SM stream a weight image from L2 into shared memory with 1-D bulk async copies, as a
// function of the copy size and of the number of copies in flight - alone and with all 148 SMs doing the same
// (the decoder's situation: every CTA streams the same 2.5 MB image, 248 KB per articulator and tile).
// Prints clk per 32 KB and B/clk/SM.  The image (2.5 MB) is L2 resident after the first pass.
// nvcc -gencode arch=compute_100a,code=sm_100a -I../../artspeech_b200/csrc -I../../include
#include <cstdio>
#include "asb_common.cuh"
using namespace asb;

constexpr int kRing = 192 * 1024;

__global__ void __launch_bounds__(32) stream_kernel(const unsigned char* image, size_t image_bytes, int piece, int depth,
                                                     int pieces, long long* out, int same_address) {
  extern __shared__ __align__(1024) unsigned char ring[];
  __shared__ uint64_t full[16];
  if (threadIdx.x == 0) {
    for (int i = 0; i < 16; ++i) mbar_init(&full[i], 1);
    fence_mbar_init();
  }
  __syncwarp();
  if (threadIdx.x != 0) return;
  // every CTA walks the image from the same offset (the decoder's CTAs are roughly in step) or from its own
  size_t off = same_address ? 0 : (size_t(blockIdx.x) * 65536) % image_bytes;
  auto issue = [&](int i) {
    const int s = i % depth;
    mbar_arrive_expect_tx(&full[s], piece);
    bulk_g2s(ring + size_t(s) * piece, image + off, piece, &full[s]);
    off += piece;
    if (off + piece > image_bytes) off = 0;
  };
  for (int i = 0; i < depth; ++i) issue(i);
  const long long t0 = clock64();
  for (int i = 0; i < pieces; ++i) {
    const int s = i % depth;
    mbar_wait(&full[s], (i / depth) & 1);
    if (i + depth < pieces) issue(i + depth);   // the slot is refilled at once (a consumer that takes no time)
  }
  out[blockIdx.x] = clock64() - t0;
}

int main() {
  const size_t image_bytes = 2560 * 1024;
  unsigned char* image;
  long long* out;
  cudaMalloc(&image, image_bytes);
  cudaMemset(image, 1, image_bytes);
  cudaMalloc(&out, 148 * sizeof(long long));
  cudaFuncSetAttribute(stream_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kRing);
  for (int same : {1, 0})
    for (int grid : {1, 148})
      for (int piece : {8192, 16384, 32768})
        for (int depth : {1, 2, 4, 6}) {
          if (piece * depth > kRing) continue;
          const int pieces = (4 << 20) / piece;   // 4 MB per CTA
          stream_kernel<<<grid, 32, kRing>>>(image, image_bytes, piece, depth, pieces, out, same);   // warm
          stream_kernel<<<grid, 32, kRing>>>(image, image_bytes, piece, depth, pieces, out, same);
          long long h[148];
          cudaError_t e = cudaMemcpy(h, out, grid * sizeof(long long), cudaMemcpyDeviceToHost);
          long long worst = 0;
          double mean = 0;
          for (int i = 0; i < grid; ++i) {
            worst = h[i] > worst ? h[i] : worst;
            mean += double(h[i]) / grid;
          }
          const double bytes = double(pieces) * piece;
          printf("%s grid %3d piece %2d KB depth %d: %7.0f clk per 32 KB (mean), %6.1f B/clk/SM mean, %6.1f worst  %s\n",
                 same ? "same-offset" : "own-offset ", grid, piece / 1024, depth, mean / (bytes / 32768.0), bytes / mean,
                 bytes / double(worst), cudaGetErrorString(e));
        }
  return 0;
}
