// Content hashes of shared-state tensors on the device.
//
// simplehash.  The hash VALUE is defined by the reference kernels (ccoip/src/cuda/simplehash_cuda.cu:
// big_reduce_kernel :95-129, final_reduce_kernel :140-156, host tail handling :234-263): a non-associative
// combine applied in the order given by a launch of min(ceil(n_vec/256), 960) blocks x 256 threads
// over contiguous slabs, shuffle trees, a 256-thread final pass, then tail words and bytes on the
// host.  Peers compare these values over the wire, so this kernel keeps that order -- but does
// it in ONE launch: the last block to finish (atomic ticket) folds the partials and the tails
// in-kernel.  No allocation, no second launch, no host tail copies, no device-wide synchronise.
// Loads are issued eight 16-byte packets ahead of the dependent combine chain.
//
// crc32c.  The reference has no device CRC (ccoip/src/cpp/ccoip_client_handler.cpp:602-605 rejects it).
// CRC is linear over GF(2): every lane folds interleaved 16-byte packets with a Horner step
// (multiply by x^4096 through a lookup table) so that warp loads stay perfectly coalesced, lanes
// are aligned with one polynomial multiplication each and XOR-reduced, and the last block
// combines the per-warp values with x^(8*bytes_after) factors.
#include <cstdio>
#include <mutex>

#include "../crc_gf.hpp"
#include "common.cuh"
#include "hash_mset.cuh"
#include "pccl_kernels.h"

namespace pccl::cuda {

constexpr uint32_t kHashThreads = 256;
constexpr uint32_t kHashMaxBlocks = 960;
// largest input with a defined hash: the packet count must fit 32 bits (2^32 - 1 packets, then up to 15 tail bytes)
constexpr uint64_t kSimpleHashMaxBytes = ((uint64_t(1) << 32) - 1) * 16 + 15;
constexpr size_t kSimpleScratchBytes = 4096; // 960 partials + ticket
constexpr uint32_t kCrcWarpsPerBlock = 8;
constexpr uint32_t kCrcMaxBlocks = 1184; // 8 per SM on 148 SMs
constexpr size_t kCrcScratchBytes = (kCrcMaxBlocks * kCrcWarpsPerBlock + 16) * sizeof(uint32_t);
constexpr size_t kScratchBytes = 64 * 1024;
static_assert(kSimpleScratchBytes + kCrcScratchBytes <= kScratchBytes, "scratch layout");

__device__ __forceinline__ uint32_t hash_combine(uint32_t a, uint32_t b) {
    a ^= b + 0x9e3779b1u;
    a = __funnelshift_l(a, a, 7);
    a *= 0x85ebca6bu;
    return a;
}

__device__ __forceinline__ uint32_t warp_fold(uint32_t v) {
#pragma unroll
    for (int offset = 16; offset > 0; offset >>= 1) v = hash_combine(v, __shfl_down_sync(0xffffffffu, v, offset));
    return v;
}

// 8 warp trees, then one tree over the warp results padded with zeros (lane >= 8 contributes 0)
__device__ uint32_t block_fold(uint32_t v) {
    __shared__ uint32_t warp_results[32];
    v = warp_fold(v);
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    __syncthreads(); // protects warp_results against the previous use
    if (lane == 0) warp_results[warp] = v;
    __syncthreads();
    uint32_t out = 0;
    if (threadIdx.x < 32) {
        out = lane < kHashThreads / 32 ? warp_results[lane] : 0;
        out = warp_fold(out);
    }
    return out;
}

__global__ void __launch_bounds__(kHashThreads) simplehash_kernel(const uint8_t *__restrict__ data, size_t n_bytes, uint32_t *scratch,
                                                                   uint64_t *out) {
    const size_t n_words = n_bytes / 4;
    const uint32_t total_vec = static_cast<uint32_t>(n_words / 4);
    uint32_t *partial = scratch;
    uint32_t *ticket = scratch + kHashMaxBlocks;
    __shared__ bool s_last;
    uint32_t result = 0;

    if (total_vec > 0) {
        const uint4 *d4 = reinterpret_cast<const uint4 *>(data);
        // slab bounds and the running index are 64-bit: with up to 2^32 - 1 packets, `total_vec + gridDim.x - 1` and
        // `i + 8 * kHashThreads` would wrap in 32 bits near the top of the range (the values themselves are unchanged)
        const uint64_t per_block = (uint64_t(total_vec) + gridDim.x - 1) / gridDim.x;
        const uint64_t begin = min(uint64_t(blockIdx.x) * per_block, uint64_t(total_vec));
        const uint64_t end = min(begin + per_block, uint64_t(total_vec));
        uint32_t h = 0;
        uint64_t i = begin + threadIdx.x;
        // eight independent loads in flight, then the serial chain
        for (; i < end && end - i > 7 * kHashThreads; i += 8 * kHashThreads) {
            uint4 v[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) v[k] = ld_stream_16(d4 + i + k * kHashThreads);
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                h = hash_combine(h, v[k].x);
                h = hash_combine(h, v[k].y);
                h = hash_combine(h, v[k].z);
                h = hash_combine(h, v[k].w);
            }
        }
        for (; i < end; i += kHashThreads) {
            const uint4 v = ld_stream_16(d4 + i);
            h = hash_combine(h, v.x);
            h = hash_combine(h, v.y);
            h = hash_combine(h, v.z);
            h = hash_combine(h, v.w);
        }
        h = block_fold(h);
        if (threadIdx.x == 0) {
            partial[blockIdx.x] = h;
            __threadfence();
            s_last = atomicAdd(ticket, 1u) == gridDim.x - 1;
        }
        __syncthreads();
        if (!s_last) return;
        __threadfence();
        uint32_t f = 0;
        for (uint32_t k = threadIdx.x; k < gridDim.x; k += kHashThreads) f = hash_combine(f, __ldcg(partial + k));
        result = block_fold(f);
    }
    if (threadIdx.x == 0) {
        const uint32_t *words = reinterpret_cast<const uint32_t *>(data);
        for (size_t w = size_t(total_vec) * 4; w < n_words; ++w) result = hash_combine(result, words[w]);
        if (const size_t tail = n_bytes % 4) {
            uint32_t last = 0;
            for (size_t k = 0; k < tail; ++k) last |= static_cast<uint32_t>(data[n_words * 4 + k]) << (8 * k);
            result = hash_combine(result, last);
        }
        *out = result;
        *ticket = 0; // leave the scratch ready for the next call
    }
}

// ------------------------------------------------------------------------------------------ mset64
// Stand-alone pass for the order-independent hash (csrc/kernels/mset_formula.hpp): used for the first hash of a tensor;
// afterwards the collective kernels keep the digest up to date as a by-product of writing the tensor.
__global__ void __launch_bounds__(512) mset64_kernel(const uint8_t *__restrict__ data, size_t n_bytes, unsigned long long *out) {
    const uint64_t n_pack = (n_bytes + 15) / 16, full = n_bytes / 16;
    const uint4 *d4 = reinterpret_cast<const uint4 *>(data);
    MsetAcc acc;
    const uint64_t stride = uint64_t(gridDim.x) * blockDim.x;
    uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    for (; i + 3 * stride < full; i += 4 * stride) { // four independent loads in flight
        uint4 v[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) v[k] = ld_stream_16(d4 + i + k * stride);
#pragma unroll
        for (int k = 0; k < 4; ++k) acc.sum += mset::term(i + k * stride, v[k].x, v[k].y, v[k].z, v[k].w);
    }
    for (; i < n_pack; i += stride) {
        if (i < full) {
            const uint4 v = ld_stream_16(d4 + i);
            acc.sum += mset::term(i, v.x, v.y, v.z, v.w);
        } else { // the partial last packet, byte by byte
            uint32_t w[4] = {0, 0, 0, 0};
            for (uint64_t b = i * 16; b < n_bytes; ++b) w[(b - i * 16) / 4] |= static_cast<uint32_t>(data[b]) << (8 * ((b - i * 16) % 4));
            acc.sum += mset::term(i, w[0], w[1], w[2], w[3]);
        }
    }
    acc.flush(out);
}

// ------------------------------------------------------------------------------------------ CRC32C

struct CrcTables {
    uint32_t byte_tab[16][256]; // byte_tab[j][b] = L(16-byte packet that is zero except byte j = b)
    uint32_t horner[4][256];    // horner[k][b]   = (b << 8k) * x^4096
};
__device__ CrcTables g_crc_tables;

__global__ void __launch_bounds__(kCrcWarpsPerBlock * 32) crc32c_kernel(const uint8_t *__restrict__ data, size_t n_bytes, uint32_t *scratch,
                                                                        uint64_t *out) {
    extern __shared__ uint32_t smem[];
    uint32_t(*byte_tab)[256] = reinterpret_cast<uint32_t(*)[256]>(smem);
    uint32_t(*horner)[256] = reinterpret_cast<uint32_t(*)[256]>(smem + 16 * 256);
    for (uint32_t i = threadIdx.x; i < 20 * 256; i += blockDim.x) smem[i] = reinterpret_cast<const uint32_t *>(&g_crc_tables)[i];
    __syncthreads();

    uint32_t *partial = scratch;
    uint32_t *ticket = scratch + kCrcMaxBlocks * kCrcWarpsPerBlock;
    const uint32_t lane = threadIdx.x & 31, warp_in_block = threadIdx.x >> 5;
    const uint32_t n_warps = gridDim.x * kCrcWarpsPerBlock, warp = blockIdx.x * kCrcWarpsPerBlock + warp_in_block;
    // bulk = whole 512-byte rows; every warp owns a contiguous span of rows
    const uint64_t rows = n_bytes / 512;
    const uint64_t rows_per_warp = (rows + n_warps - 1) / n_warps;
    const uint64_t row_lo = min(uint64_t(warp) * rows_per_warp, rows), row_hi = min(row_lo + rows_per_warp, rows);

    uint32_t acc = 0;
    const uint4 *d4 = reinterpret_cast<const uint4 *>(data);
    auto fold_packet = [&](uint32_t a, const uint4 v) {
        // A' = A * x^4096 + L(packet): four lookups for the Horner step, sixteen for the bytes
        uint32_t r = horner[0][a & 0xff] ^ horner[1][(a >> 8) & 0xff] ^ horner[2][(a >> 16) & 0xff] ^ horner[3][a >> 24];
        const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            r ^= byte_tab[4 * j + 0][w[j] & 0xff];
            r ^= byte_tab[4 * j + 1][(w[j] >> 8) & 0xff];
            r ^= byte_tab[4 * j + 2][(w[j] >> 16) & 0xff];
            r ^= byte_tab[4 * j + 3][w[j] >> 24];
        }
        return r;
    };
    uint64_t row = row_lo;
    for (; row + 4 <= row_hi; row += 4) {
        uint4 v[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) v[k] = ld_stream_16(d4 + (row + k) * 32 + lane);
#pragma unroll
        for (int k = 0; k < 4; ++k) acc = fold_packet(acc, v[k]);
    }
    for (; row < row_hi; ++row) acc = fold_packet(acc, ld_stream_16(d4 + row * 32 + lane));
    // align the lanes: lane l is followed by (31-l) packets inside its row
    acc = gf::mul(acc, gf::x_pow(uint64_t(31 - lane) * 128));
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc ^= __shfl_xor_sync(0xffffffffu, acc, o);

    __shared__ bool s_last;
    if (lane == 0) partial[warp] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        s_last = atomicAdd(ticket, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    // combine: the value of warp w is followed by all bulk bytes after its span.  All spans but the
    // last have the same length, so a thread walks a contiguous run of warps with a Horner step
    // (one multiplication by X = x^(8*span) per warp) and pays a single x^n computation for the
    // bytes that follow its run.
    uint32_t total = 0;
    {
        const uint32_t per_thread = (n_warps + blockDim.x - 1) / blockDim.x;
        const uint32_t w_lo = min(threadIdx.x * per_thread, n_warps), w_hi = min(w_lo + per_thread, n_warps);
        const uint32_t X = gf::x_pow(rows_per_warp * 512 * 8);
        uint32_t run = 0;
        uint64_t last_row_end = 0;
        for (uint32_t w = w_lo; w < w_hi; ++w) {
            const uint64_t lo = min(uint64_t(w) * rows_per_warp, rows), hi = min(lo + rows_per_warp, rows);
            if (hi == lo) break; // spans are assigned in order: the rest is empty
            // shift what we have by the length of THIS span, then append it
            run = hi - lo == rows_per_warp ? gf::mul(run, X) : gf::mul(run, gf::x_pow((hi - lo) * 512 * 8));
            run ^= __ldcg(partial + w);
            last_row_end = hi;
        }
        if (last_row_end) total = gf::mul(run, gf::x_pow((rows - last_row_end) * 512 * 8));
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) total ^= __shfl_xor_sync(0xffffffffu, total, o);
    __shared__ uint32_t s_warp_totals[kCrcWarpsPerBlock];
    if (lane == 0) s_warp_totals[warp_in_block] = total;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t state = 0;
        for (uint32_t w = 0; w < kCrcWarpsPerBlock; ++w) state ^= s_warp_totals[w];
        // the all-ones initial register, advanced over the bulk part
        state ^= gf::mul(0xFFFFFFFFu, gf::x_pow(rows * 512 * 8));
        // tail (< 512 bytes) byte by byte; byte_tab[15] is the plain one-byte CRC table
        // advanced by zero bytes, i.e. T0
        for (uint64_t i = rows * 512; i < n_bytes; ++i) state = (state >> 8) ^ byte_tab[15][(state ^ data[i]) & 0xff];
        *out = static_cast<uint32_t>(~state);
        *ticket = 0;
    }
}

static void build_crc_tables(CrcTables &t) {
    // L(packet with single byte b at position j) = T0[b] advanced over the (15-j) bytes that follow
    uint32_t t0[256];
    for (uint32_t b = 0; b < 256; ++b) {
        uint32_t c = b;
        for (int k = 0; k < 8; ++k) c = (c >> 1) ^ ((c & 1u) ? gf::kCrc32cPoly : 0u);
        t0[b] = c;
    }
    for (int j = 0; j < 16; ++j) {
        const uint32_t factor = gf::x_pow(uint64_t(15 - j) * 8);
        for (uint32_t b = 0; b < 256; ++b) t.byte_tab[j][b] = gf::mul(t0[b], factor);
    }
    const uint32_t x4096 = gf::x_pow(4096);
    for (int k = 0; k < 4; ++k)
        for (uint32_t b = 0; b < 256; ++b) t.horner[k][b] = gf::mul(b << (8 * k), x4096);
}

static int ensure_crc_tables() {
    static std::mutex mu;
    static bool done[64] = {false};
    int dev = 0;
    cudaError_t rc = cudaGetDevice(&dev);
    if (rc != cudaSuccess) return rc;
    std::lock_guard<std::mutex> lock(mu);
    if (dev < 64 && done[dev]) return 0;
    static CrcTables host;
    static bool built = false;
    if (!built) {
        build_crc_tables(host);
        built = true;
    }
    rc = cudaMemcpyToSymbol(g_crc_tables, &host, sizeof host);
    if (rc != cudaSuccess) return rc;
    rc = cudaFuncSetAttribute(crc32c_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 20 * 1024);
    if (rc != cudaSuccess) return rc;
    if (dev < 64) done[dev] = true;
    return 0;
}

} // namespace pccl::cuda

using namespace pccl::cuda;

extern "C" size_t pcclKernelHashScratchBytes(void) { return kScratchBytes; }

extern "C" int pcclKernelSimpleHash(const void *data, size_t n_bytes, void *scratch, uint64_t *out_device, void *stream) {
    if (!scratch || !out_device) return cudaErrorInvalidValue;
    if (n_bytes == 0) return cudaMemsetAsync(out_device, 0, sizeof(uint64_t), static_cast<cudaStream_t>(stream));
    if (reinterpret_cast<uintptr_t>(data) % 16 != 0) return cudaErrorMisalignedAddress;
    // the hash counts 16-byte packets in 32 bits (so does the reference kernel): larger inputs have no defined value
    if (static_cast<uint64_t>(n_bytes) > kSimpleHashMaxBytes) return cudaErrorInvalidValue;
    const size_t n_vec = n_bytes / 16;
    uint32_t grid = static_cast<uint32_t>(std::min<size_t>((n_vec + kHashThreads - 1) / kHashThreads, kHashMaxBlocks));
    if (grid == 0) grid = 1;
    simplehash_kernel<<<grid, kHashThreads, 0, static_cast<cudaStream_t>(stream)>>>(static_cast<const uint8_t *>(data), n_bytes,
                                                                                   static_cast<uint32_t *>(scratch), out_device);
    return cudaGetLastError();
}

extern "C" int pcclKernelMset64Raw(const void *data, size_t n_bytes, uint64_t *out_device, void *stream_v) {
    if (!out_device || (!data && n_bytes > 0)) return cudaErrorInvalidValue;
    const auto stream = static_cast<cudaStream_t>(stream_v);
    cudaError_t rc = cudaMemsetAsync(out_device, 0, sizeof(uint64_t), stream);
    if (rc != cudaSuccess || n_bytes == 0) return rc;
    if (reinterpret_cast<uintptr_t>(data) % 16 != 0) return cudaErrorMisalignedAddress;
    const uint64_t n_pack = (n_bytes + 15) / 16;
    const int grid = static_cast<int>(std::min<uint64_t>(148 * 4, (n_pack + 2047) / 2048));
    mset64_kernel<<<std::max(1, grid), 512, 0, stream>>>(static_cast<const uint8_t *>(data), n_bytes, reinterpret_cast<unsigned long long *>(out_device));
    return cudaGetLastError();
}

extern "C" uint64_t pcclExtMsetFinalize(uint64_t raw_sum, uint64_t n_bytes) { return pccl::mset::finalize(raw_sum, n_bytes); }

extern "C" int pcclKernelCrc32c(const void *data, size_t n_bytes, void *scratch, uint64_t *out_device, void *stream) {
    if (!scratch || !out_device) return cudaErrorInvalidValue;
    if (reinterpret_cast<uintptr_t>(data) % 16 != 0 && n_bytes >= 512) return cudaErrorMisalignedAddress;
    if (const int rc = ensure_crc_tables()) return rc;
    const uint64_t rows = n_bytes / 512;
    // enough warps that each gets at least ~16 rows, at most the persistent grid
    uint32_t grid = static_cast<uint32_t>(std::min<uint64_t>(kCrcMaxBlocks, (rows / 16 + kCrcWarpsPerBlock - 1) / kCrcWarpsPerBlock));
    if (grid == 0) grid = 1;
    auto *crc_scratch = reinterpret_cast<uint32_t *>(static_cast<uint8_t *>(scratch) + kSimpleScratchBytes);
    crc32c_kernel<<<grid, kCrcWarpsPerBlock * 32, 20 * 1024, static_cast<cudaStream_t>(stream)>>>(
        static_cast<const uint8_t *>(data), n_bytes, crc_scratch, out_device);
    return cudaGetLastError();
}

// Per-device scratch + stream kept by the library for the synchronous convenience entry.
namespace {
struct HashContext {
    void *scratch = nullptr;
    uint64_t *out = nullptr;
    cudaStream_t stream = nullptr;
    cudaEvent_t ready = nullptr; // orders the hash after work queued on the legacy default stream
    uint64_t *host_out = nullptr; // pinned landing place of the result
};
HashContext *hash_context(int &rc) {
    static std::mutex mu;
    static HashContext ctx[64];
    int dev = 0;
    rc = cudaGetDevice(&dev);
    if (rc != cudaSuccess || dev >= 64) {
        if (rc == cudaSuccess) rc = cudaErrorInvalidDevice;
        return nullptr;
    }
    std::lock_guard<std::mutex> lock(mu);
    HashContext &c = ctx[dev];
    if (!c.scratch) {
        if ((rc = cudaMalloc(&c.scratch, kScratchBytes)) != cudaSuccess) return nullptr;
        if ((rc = cudaMemset(c.scratch, 0, kScratchBytes)) != cudaSuccess) return nullptr;
        if ((rc = cudaMalloc(&c.out, sizeof(uint64_t))) != cudaSuccess) return nullptr;
        if ((rc = cudaStreamCreateWithFlags(&c.stream, cudaStreamNonBlocking)) != cudaSuccess) return nullptr;
        if ((rc = cudaEventCreateWithFlags(&c.ready, cudaEventDisableTiming)) != cudaSuccess) return nullptr;
        if ((rc = cudaHostAlloc(reinterpret_cast<void **>(&c.host_out), sizeof(uint64_t), cudaHostAllocDefault)) != cudaSuccess) return nullptr;
    }
    return &c;
}
std::mutex g_hash_call_mu; // the scratch is shared: one hash at a time per process
} // namespace

extern "C" int pcclKernelHashSync(const void *data, size_t n_bytes, int hash_type, uint64_t *out_host) {
    if (!out_host) return cudaErrorInvalidValue;
    int rc = 0;
    HashContext *c = hash_context(rc);
    if (!c) return rc;
    std::lock_guard<std::mutex> lock(g_hash_call_mu);
    // the data may have been produced on the legacy default stream
    cudaEventRecord(c->ready, cudaStreamLegacy);
    cudaStreamWaitEvent(c->stream, c->ready, 0);
    rc = hash_type == 0   ? pcclKernelSimpleHash(data, n_bytes, c->scratch, c->out, c->stream)
         : hash_type == 1 ? pcclKernelCrc32c(data, n_bytes, c->scratch, c->out, c->stream)
                          : pcclKernelMset64Raw(data, n_bytes, c->out, c->stream);
    if (rc != cudaSuccess) return rc;
    if ((rc = cudaMemcpyAsync(c->host_out, c->out, sizeof(uint64_t), cudaMemcpyDeviceToHost, c->stream)) != cudaSuccess) return rc;
    if ((rc = cudaStreamSynchronize(c->stream)) != cudaSuccess) return rc;
    *out_host = hash_type == 2 ? pccl::mset::finalize(*c->host_out, n_bytes) : *c->host_out;
    return 0;
}

// Many buffers, one wait: the kernels run back to back on the library's hash stream (they share the one scratch area, which
// every kernel leaves ready for the next), each writes its own result word, one copy brings all results back.
extern "C" int pcclKernelHashManySync(const void *const *data, const size_t *n_bytes, size_t count, int hash_type, uint64_t *out_host) {
    if (!data || !n_bytes || !out_host) return cudaErrorInvalidValue;
    if (count == 0) return 0;
    int rc = 0;
    HashContext *c = hash_context(rc);
    if (!c) return rc;
    std::lock_guard<std::mutex> lock(g_hash_call_mu);
    static thread_local uint64_t *dev_out = nullptr, *host_out = nullptr;
    static thread_local size_t capacity = 0;
    static thread_local int owner_device = -1;
    int dev = 0;
    cudaGetDevice(&dev);
    if (capacity < count || owner_device != dev) {
        if (dev_out) cudaFree(dev_out);
        if (host_out) cudaFreeHost(host_out);
        dev_out = host_out = nullptr;
        capacity = 0;
        const size_t want = std::max<size_t>(count, 1024);
        if ((rc = cudaMalloc(reinterpret_cast<void **>(&dev_out), want * sizeof(uint64_t))) != cudaSuccess) return rc;
        if ((rc = cudaHostAlloc(reinterpret_cast<void **>(&host_out), want * sizeof(uint64_t), cudaHostAllocDefault)) != cudaSuccess) return rc;
        capacity = want;
        owner_device = dev;
    }
    cudaEventRecord(c->ready, cudaStreamLegacy);
    cudaStreamWaitEvent(c->stream, c->ready, 0);
    for (size_t i = 0; i < count; ++i) {
        rc = hash_type == 0   ? pcclKernelSimpleHash(data[i], n_bytes[i], c->scratch, dev_out + i, c->stream)
             : hash_type == 1 ? pcclKernelCrc32c(data[i], n_bytes[i], c->scratch, dev_out + i, c->stream)
                              : pcclKernelMset64Raw(data[i], n_bytes[i], dev_out + i, c->stream);
        if (rc != cudaSuccess) return rc;
    }
    if ((rc = cudaMemcpyAsync(host_out, dev_out, count * sizeof(uint64_t), cudaMemcpyDeviceToHost, c->stream)) != cudaSuccess) return rc;
    if ((rc = cudaStreamSynchronize(c->stream)) != cudaSuccess) return rc;
    for (size_t i = 0; i < count; ++i) out_host[i] = hash_type == 2 ? pccl::mset::finalize(host_out[i], n_bytes[i]) : host_out[i];
    return 0;
}

// ---------------------------------------------------------------------------------------------
// pcclKernelSimpleHashNoAlloc -- the same hash for callers that must not allocate.
//
// Who that is: code that runs while a CUDA stream capture is open on the thread, code that runs
// when the framework's caching allocator already owns all device memory (a fresh cudaMalloc would
// fail or force the allocator to release and synchronise), and latency-sensitive paths where the
// device-wide synchronisation inside cudaMalloc/cudaFree would stall every other stream.  The
// regular entry points above either take a caller-provided scratch buffer or create one lazily
// with cudaMalloc; this one owns its scratch as a static device array instead, so every GPU that
// loads libpccl.so gives up kNoAllocScratchBytes until the process exits.
//
// The scratch holds one 32-bit partial per thread block (at most 960) and the ticket counter the
// last block uses to detect that it is last.  The hash is defined for inputs of up to 2^32 - 1
// sixteen-byte packets (the kernel, like the reference, counts packets in 32 bits), so that is the
// size limit; anything larger is refused.
//
// Rules for callers:
//   * the first call on each device looks up the device address of the scratch array
//     (cudaGetSymbolAddress).  Make that first call before any stream capture begins.
//   * all calls on one device share the one scratch array.  The ticket counter is cleared on the
//     stream before every launch (a failed earlier launch could have left it dirty), the partials
//     need no clearing because every block overwrites its own before the fold.  Two calls whose
//     kernels overlap on the GPU (different streams, no ordering) race on the array and return
//     garbage: order them.
//   * no cudaMalloc, cudaFree or other allocating call is made, on any path.
// ---------------------------------------------------------------------------------------------
namespace pccl::cuda {

constexpr size_t kNoAllocScratchBytes = 4096; // 960 partials + ticket, rounded up
constexpr uint64_t kNoAllocMaxBytes = kSimpleHashMaxBytes; // 2^32 - 1 packets plus a tail of up to 15 bytes
static_assert((kHashMaxBlocks + 1) * sizeof(uint32_t) <= kNoAllocScratchBytes, "static scratch too small");

static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocScratchBytes];

// In host code the array's name is only a handle for the runtime, not a device address; the real
// address is fetched once per device with that device current.
static int noalloc_scratch_address(uint32_t **out) {
    static std::mutex mu;
    static uint32_t *address[64] = {nullptr};
    int dev = 0;
    cudaError_t rc = cudaGetDevice(&dev);
    if (rc != cudaSuccess) return rc;
    if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lock(mu);
    if (!address[dev]) {
        void *p = nullptr;
        rc = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (rc != cudaSuccess) return rc;
        address[dev] = static_cast<uint32_t *>(p);
    }
    *out = address[dev];
    return 0;
}

} // namespace pccl::cuda

extern "C" PCCL_EXPORT size_t pcclKernelSimpleHashNoAllocMaxBytes(void) { return static_cast<size_t>(kNoAllocMaxBytes); }

extern "C" PCCL_EXPORT int pcclKernelSimpleHashNoAlloc(const void *data, size_t n_bytes, uint64_t *out_device, void *stream_v) {
    if (!out_device || (!data && n_bytes > 0)) return cudaErrorInvalidValue;
    if (static_cast<uint64_t>(n_bytes) > kNoAllocMaxBytes) {
        fprintf(stderr, "pcclKernelSimpleHashNoAlloc: %zu bytes exceed the supported maximum of %llu bytes\n", n_bytes,
                static_cast<unsigned long long>(kNoAllocMaxBytes));
        return cudaErrorInvalidValue;
    }
    const auto stream = static_cast<cudaStream_t>(stream_v);
    if (n_bytes == 0) return cudaMemsetAsync(out_device, 0, sizeof(uint64_t), stream);
    if (reinterpret_cast<uintptr_t>(data) % 16 != 0) return cudaErrorMisalignedAddress;
    uint32_t *scratch = nullptr;
    if (const int rc = noalloc_scratch_address(&scratch)) return rc;
    // the ticket must be zero when the kernel starts
    cudaError_t rc = cudaMemsetAsync(scratch + kHashMaxBlocks, 0, sizeof(uint32_t), stream);
    if (rc != cudaSuccess) return rc;
    const size_t n_vec = n_bytes / 16;
    uint32_t grid = static_cast<uint32_t>(std::min<size_t>((n_vec + kHashThreads - 1) / kHashThreads, kHashMaxBlocks));
    if (grid == 0) grid = 1;
    simplehash_kernel<<<grid, kHashThreads, 0, stream>>>(static_cast<const uint8_t *>(data), n_bytes, scratch, out_device);
    return cudaGetLastError();
}
