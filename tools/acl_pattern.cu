// acl_pattern.cu — what would an ACL structure with a given memory access PATTERN cost next to the packet stream?
// Not a classifier: each "packet" reads its 64-byte record, then performs `wide` independent loads (phase 0 of an
// equivalence-class scheme: one small table per key chunk) followed by `deep` dependent loads (combine steps, or the
// levels of a decision tree when wide = 0), all in L2-resident tables indexed by mixed-up bits of the previous value,
// and writes 8 bytes.  Compare the patterns  tree: wide 0, deep 7   and   classes: wide 6, deep 3 (3 loads per level).
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/_build/acl_pattern tools/acl_pattern.cu
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>

__device__ __forceinline__ uint32_t mix(uint32_t x) {
    x ^= x >> 16;
    x *= 0x7feb352du;
    x ^= x >> 15;
    x *= 0x846ca68bu;
    return x ^ (x >> 16);
}

template <int WIDE, int DEEP, int PER_LEVEL>
__global__ void __launch_bounds__(256, 4) pattern(const uint4* __restrict__ hdr, uint64_t* __restrict__ out, size_t n,
                                                  const uint16_t* __restrict__ small, uint32_t small_mask,
                                                  const uint32_t* __restrict__ big, uint32_t big_mask) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const uint4 a = hdr[4 * i], b = hdr[4 * i + 1], c = hdr[4 * i + 2], d = hdr[4 * i + 3];
        uint32_t key[8] = {a.x ^ c.x, a.y ^ c.y, a.z ^ c.z, a.w ^ c.w, b.x ^ d.x, b.y ^ d.y, b.z ^ d.z, b.w ^ d.w};
        uint32_t v = key[0];
#pragma unroll
        for (int w = 0; w < WIDE; ++w)  // independent: one 2-byte class id per chunk, tables 128 KB apart
            v += __ldg(small + ((size_t)w << 16) + (mix(key[w & 7] + w) & small_mask & 0xFFFFu));
#pragma unroll
        for (int l = 0; l < DEEP; ++l) {  // dependent levels, PER_LEVEL loads each
            uint32_t nv = 0;
#pragma unroll
            for (int p = 0; p < PER_LEVEL; ++p) nv += __ldg(big + (mix(v + 0x9E37u * p + l) & big_mask));
            v = nv;
        }
        out[i] = v;
    }
}

template <int WIDE, int DEEP, int PER_LEVEL>
float run(const uint4* hdr, uint64_t* out, size_t n, const uint16_t* small, const uint32_t* big, uint32_t big_mask, int reps) {
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    pattern<WIDE, DEEP, PER_LEVEL><<<148 * 8, 256>>>(hdr, out, n, small, 0xFFFFu, big, big_mask);
    cudaEventRecord(e0);
    for (int r = 0; r < reps; ++r) pattern<WIDE, DEEP, PER_LEVEL><<<148 * 8, 256>>>(hdr, out, n, small, 0xFFFFu, big, big_mask);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    return ms * 1e3f / reps / (float)(n >> 20);  // us per Mi packets
}

int main() {
    const size_t n = 8u << 20;  // 8 Mi packets: 512 MB of records, far beyond L2
    uint4* hdr;
    uint64_t* out;
    uint16_t* small;
    uint32_t* big;
    cudaMalloc(&hdr, n * 64);
    cudaMalloc(&out, n * 8);
    cudaMalloc(&small, 8u << 17);  // 8 chunk tables of 64 Ki u16
    std::vector<uint32_t> h(n * 16 / 64);  // random-ish fill, repeated
    for (size_t i = 0; i < h.size(); ++i) h[i] = (uint32_t)(i * 2654435761u) ^ (uint32_t)(i >> 7);
    for (size_t off = 0; off < n * 64; off += h.size() * 4) cudaMemcpy((char*)hdr + off, h.data(), h.size() * 4, cudaMemcpyHostToDevice);
    cudaMemset(small, 1, 8u << 17);
    for (int mb : {4, 16}) {
        const uint32_t words = (uint32_t)mb << 18;
        cudaMalloc(&big, (size_t)words * 4);
        cudaMemset(big, 3, (size_t)words * 4);
        const uint32_t m = words - 1;
        printf("tables of %d MB (L2-resident), us per Mi packets (stream floor ~11.9):\n", mb);
        printf("  stream only                         : %6.2f\n", run<0, 0, 1>(hdr, out, n, small, big, m, 10));
        printf("  tree      : 4 dependent loads       : %6.2f\n", run<0, 4, 1>(hdr, out, n, small, big, m, 10));
        printf("  tree      : 7 dependent loads       : %6.2f\n", run<0, 7, 1>(hdr, out, n, small, big, m, 10));
        printf("  tree      : 10 dependent loads      : %6.2f\n", run<0, 10, 1>(hdr, out, n, small, big, m, 10));
        printf("  classes   : 6 wide + 3 levels x 1   : %6.2f\n", run<6, 3, 1>(hdr, out, n, small, big, m, 10));
        printf("  classes   : 6 wide + 3 levels x 2   : %6.2f\n", run<6, 3, 2>(hdr, out, n, small, big, m, 10));
        printf("  classes   : 8 wide + 3 levels x 3   : %6.2f\n", run<8, 3, 3>(hdr, out, n, small, big, m, 10));
        printf("  classes   : 4 wide + 2 levels x 2   : %6.2f\n", run<4, 2, 2>(hdr, out, n, small, big, m, 10));
        cudaFree(big);
    }
    return 0;
}
