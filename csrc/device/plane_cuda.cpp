// CUDA implementation of the device plane (see plane.hpp for the design).
//
// Driver-API entry points (virtual memory management, multicast) are resolved at run time through
// cudaGetDriverEntryPoint, so libpccl.so has no link-time dependency on libcuda and loads on
// machines without a GPU.
#ifndef PCCL_NO_CUDA
#include <cuda.h>
#include <cuda_runtime.h>
#include <poll.h>
#include <sys/socket.h>
#include <sys/un.h>
#include <unistd.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <cstring>
#include <deque>
#include <map>
#include <mutex>
#include <thread>
#include <unordered_map>

#include "../util/env.hpp"
#include "../util/log.hpp"
#include "pccl_kernels.h"
#include "plane.hpp"

namespace pccl::device {

namespace {

using clock_type = std::chrono::steady_clock;

// ------------------------------------------------------------------------------------------ driver API
struct Driver {
    bool ok = false;
    CUresult (*MemGetAllocationGranularity)(size_t *, const CUmemAllocationProp *, CUmemAllocationGranularity_flags) = nullptr;
    CUresult (*MemCreate)(CUmemGenericAllocationHandle *, size_t, const CUmemAllocationProp *, unsigned long long) = nullptr;
    CUresult (*MemRelease)(CUmemGenericAllocationHandle) = nullptr;
    CUresult (*MemAddressReserve)(CUdeviceptr *, size_t, size_t, CUdeviceptr, unsigned long long) = nullptr;
    CUresult (*MemAddressFree)(CUdeviceptr, size_t) = nullptr;
    CUresult (*MemMap)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long) = nullptr;
    CUresult (*MemUnmap)(CUdeviceptr, size_t) = nullptr;
    CUresult (*MemSetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc *, size_t) = nullptr;
    CUresult (*MemExportToShareableHandle)(void *, CUmemGenericAllocationHandle, CUmemAllocationHandleType, unsigned long long) = nullptr;
    CUresult (*MemImportFromShareableHandle)(CUmemGenericAllocationHandle *, void *, CUmemAllocationHandleType) = nullptr;
    CUresult (*DeviceGet)(CUdevice *, int) = nullptr;
    CUresult (*DeviceGetAttribute)(int *, CUdevice_attribute, CUdevice) = nullptr;
    // NVSwitch multicast (optional: absence only disables the NVLS path)
    bool multicast = false;
    CUresult (*MulticastCreate)(CUmemGenericAllocationHandle *, const CUmulticastObjectProp *) = nullptr;
    CUresult (*MulticastAddDevice)(CUmemGenericAllocationHandle, CUdevice) = nullptr;
    CUresult (*MulticastBindMem)(CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle, size_t, size_t, unsigned long long) = nullptr;
    CUresult (*MulticastUnbind)(CUmemGenericAllocationHandle, CUdevice, size_t, size_t) = nullptr;
    CUresult (*MulticastGetGranularity)(size_t *, const CUmulticastObjectProp *, CUmulticastGranularity_flags) = nullptr;
};

const Driver &driver() {
    static const Driver d = [] {
        Driver x;
        int count = 0;
        if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0) {
            (void) cudaGetLastError();
            return x;
        }
        bool all = true;
        auto load = [&](const char *name, auto &fn) {
            void *p = nullptr;
            cudaDriverEntryPointQueryResult q;
            if (cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess || !p) {
                (void) cudaGetLastError();
                all = false;
                return;
            }
            fn = reinterpret_cast<std::remove_reference_t<decltype(fn)>>(p);
        };
        load("cuMemGetAllocationGranularity", x.MemGetAllocationGranularity);
        load("cuMemCreate", x.MemCreate);
        load("cuMemRelease", x.MemRelease);
        load("cuMemAddressReserve", x.MemAddressReserve);
        load("cuMemAddressFree", x.MemAddressFree);
        load("cuMemMap", x.MemMap);
        load("cuMemUnmap", x.MemUnmap);
        load("cuMemSetAccess", x.MemSetAccess);
        load("cuMemExportToShareableHandle", x.MemExportToShareableHandle);
        load("cuMemImportFromShareableHandle", x.MemImportFromShareableHandle);
        load("cuDeviceGet", x.DeviceGet);
        load("cuDeviceGetAttribute", x.DeviceGetAttribute);
        x.ok = all;
        all = true;
        load("cuMulticastCreate", x.MulticastCreate);
        load("cuMulticastAddDevice", x.MulticastAddDevice);
        load("cuMulticastBindMem", x.MulticastBindMem);
        load("cuMulticastUnbind", x.MulticastUnbind);
        load("cuMulticastGetGranularity", x.MulticastGetGranularity);
        x.multicast = all;
        return x;
    }();
    return d;
}

#define CU_TRY(expr)                                                                                                   \
    do {                                                                                                               \
        const CUresult rc_ = (expr);                                                                                   \
        if (rc_ != CUDA_SUCCESS) {                                                                                     \
            PLOG(WARN) << #expr << " failed with CUresult " << static_cast<int>(rc_);                                  \
            return false;                                                                                              \
        }                                                                                                              \
    } while (0)
#define RT_TRY(expr)                                                                                                   \
    do {                                                                                                               \
        const cudaError_t rc_ = (expr);                                                                                \
        if (rc_ != cudaSuccess) {                                                                                      \
            PLOG(WARN) << #expr << " failed: " << cudaGetErrorString(rc_);                                             \
            (void) cudaGetLastError();                                                                                 \
            return false;                                                                                              \
        }                                                                                                              \
    } while (0)

// ------------------------------------------------------------------------------------------ unix rendezvous
constexpr uint32_t kMagic = 0x50434231; // "PCB1"
enum : uint32_t { kReqArena = 1, kReqMulticast = 2, kReqMulticastAdded = 3, kReqMulticastReady = 4 };

struct RendezvousRequest {
    uint32_t magic, kind;
    uint8_t requester[16];
    uint64_t ring_key; // identifies the membership epoch a multicast object belongs to
};
struct RendezvousReply {
    uint32_t magic, status; // status 0 = ok
    uint64_t arena_bytes, slot_bytes, sym_bytes;
    uint32_t n_slots, pid;
    uint8_t gpu_uuid[16];
    uint32_t multicast_ok, added_count;
};

sockaddr_un rendezvous_address(const Uuid &u, socklen_t &len) {
    sockaddr_un sa{};
    sa.sun_family = AF_UNIX;
    const std::string name = "pccl-b200-" + u.hex();
    sa.sun_path[0] = '\0'; // abstract namespace: no file system entry, vanishes with the process
    std::memcpy(sa.sun_path + 1, name.data(), name.size());
    len = static_cast<socklen_t>(offsetof(sockaddr_un, sun_path) + 1 + name.size());
    return sa;
}

bool send_with_fd(int sock, const void *data, size_t n, int fd) {
    msghdr msg{};
    iovec iov{const_cast<void *>(data), n};
    msg.msg_iov = &iov;
    msg.msg_iovlen = 1;
    alignas(cmsghdr) char ctrl[CMSG_SPACE(sizeof(int))] = {0};
    if (fd >= 0) {
        msg.msg_control = ctrl;
        msg.msg_controllen = sizeof ctrl;
        cmsghdr *c = CMSG_FIRSTHDR(&msg);
        c->cmsg_level = SOL_SOCKET;
        c->cmsg_type = SCM_RIGHTS;
        c->cmsg_len = CMSG_LEN(sizeof(int));
        std::memcpy(CMSG_DATA(c), &fd, sizeof(int));
    }
    return ::sendmsg(sock, &msg, MSG_NOSIGNAL) == static_cast<ssize_t>(n);
}

bool recv_with_fd(int sock, void *data, size_t n, int &fd, int timeout_ms) {
    fd = -1;
    pollfd p{sock, POLLIN, 0};
    if (::poll(&p, 1, timeout_ms) <= 0) return false;
    msghdr msg{};
    iovec iov{data, n};
    msg.msg_iov = &iov;
    msg.msg_iovlen = 1;
    alignas(cmsghdr) char ctrl[CMSG_SPACE(sizeof(int))] = {0};
    msg.msg_control = ctrl;
    msg.msg_controllen = sizeof ctrl;
    if (::recvmsg(sock, &msg, MSG_CMSG_CLOEXEC) != static_cast<ssize_t>(n)) return false;
    for (cmsghdr *c = CMSG_FIRSTHDR(&msg); c; c = CMSG_NXTHDR(&msg, c))
        if (c->cmsg_level == SOL_SOCKET && c->cmsg_type == SCM_RIGHTS) std::memcpy(&fd, CMSG_DATA(c), sizeof(int));
    return true;
}

size_t round_up(size_t v, size_t g) { return (v + g - 1) / g * g; }

// ------------------------------------------------------------------------------------------ the plane
class CudaPlane final : public Plane {
public:
    CudaPlane() {
        usable_ = cuda_runtime_usable() && driver().ok && env_flag("PCCL_DEVICE_PLANE", true);
        if (usable_ && cudaGetDevice(&device_) != cudaSuccess) usable_ = false;
        if (usable_) {
            cudaDeviceProp prop{};
            if (cudaGetDeviceProperties(&prop, device_) == cudaSuccess) {
                std::memcpy(gpu_uuid_, prop.uuid.bytes, 16);
                clock_khz_ = prop.clockRate > 0 ? prop.clockRate : 1900000;
            }
        }
        slot_bytes_ = round_up(env_u64("PCCL_ARENA_SLOT_MB", 768) << 20, 2u << 20);
        n_slots_ = static_cast<uint32_t>(std::clamp<uint64_t>(env_u64("PCCL_ARENA_SLOTS", 2), 1, 15));
        // symmetric heap = [caller's allocations][library scratch].  The scratch region has a FIXED offset, so peers
        // with different allocation histories (late joiners, different tensor lists in the past) still agree on where
        // an internally staged operand lives -- symmetric operands are addressed with one offset in every arena.
        sym_bytes_ = round_up(env_u64("PCCL_SYMMETRIC_MB", 2048) << 20, 2u << 20);
        internal_bytes_ = std::min<size_t>(sym_bytes_ / 2, round_up(env_u64("PCCL_INTERNAL_SCRATCH_MB", 1024) << 20, 2u << 20));
        timeout_ms_ = env_u64("PCCL_DEVICE_TIMEOUT_MS", 30000);
        max_blocks_ = static_cast<int>(env_u64("PCCL_DEVICE_MAX_BLOCKS", 0));
        nvls_min_world_ = static_cast<uint32_t>(env_u64("PCCL_NVLS_MIN_WORLD", 4));
        tune_ = env_flag("PCCL_DEVICE_TUNE", false);
        pipeline_min_bytes_ = env_u64("PCCL_PIPELINE_MIN_MB", 256) << 20; // below that the extra barriers eat the overlap
        kernel_pipeline_ = env_flag("PCCL_KERNEL_PIPELINE", true);
        pipeline_segments_ = env_u64("PCCL_PIPELINE_SEGMENTS", 0);
        pipeline_seg_bytes_ = std::max<uint64_t>(1, env_u64("PCCL_PIPELINE_SEGMENT_MB", 64)) << 20;
        nvls_min_bytes_ = env_u64("PCCL_NVLS_MIN_BYTES", 1u << 20);
    }
    ~CudaPlane() override { stop(); }

    void start(const Uuid &self) override {
        self_ = self;
        if (!usable_ || started_) return;
        listen_fd_ = ::socket(AF_UNIX, SOCK_STREAM | SOCK_CLOEXEC, 0);
        if (listen_fd_ < 0) return;
        socklen_t len;
        const sockaddr_un sa = rendezvous_address(self, len);
        if (::bind(listen_fd_, reinterpret_cast<const sockaddr *>(&sa), len) != 0 || ::listen(listen_fd_, 64) != 0) {
            ::close(listen_fd_);
            listen_fd_ = -1;
            return;
        }
        started_ = true;
        server_ = std::thread([this] { serve(); });
    }

    void stop() override {
        if (started_) {
            stopping_ = true;
            if (server_.joinable()) server_.join();
            ::close(listen_fd_);
            listen_fd_ = -1;
            started_ = false;
        }
        std::lock_guard<std::mutex> lock(mu_);
        if (!usable_) return;
        cudaSetDevice(device_);
        if (arena_ptr_) teardown_multicast();
        for (auto &[u, p] : peers_) unmap_peer(p);
        peers_.clear();
        for (auto &s : slots_) {
            if (s.stream) cudaStreamSynchronize(s.stream);
            if (s.stream) cudaStreamDestroy(s.stream);
            if (s.copy_in) cudaStreamDestroy(s.copy_in);
            if (s.copy_out) cudaStreamDestroy(s.copy_out);
            for (auto *v : {&s.ev_in, &s.ev_kernel, &s.ev_out})
                for (cudaEvent_t e : *v) cudaEventDestroy(e);
            if (s.start) cudaEventDestroy(s.start);
            if (s.stop) cudaEventDestroy(s.stop);
            if (s.host_words) cudaFreeHost(s.host_words);
            if (s.backup) cudaFree(s.backup);
        }
        slots_.clear();
        if (util_stream_) cudaStreamDestroy(util_stream_), util_stream_ = nullptr;
        if (arena_ptr_) {
            driver().MemUnmap(arena_ptr_, arena_bytes_);
            driver().MemAddressFree(arena_ptr_, arena_bytes_);
            driver().MemRelease(arena_handle_);
            arena_ptr_ = 0;
        }
    }

    void on_ring_changed(const std::vector<Uuid> &ring) override {
        std::lock_guard<std::mutex> lock(mu_);
        ring_ = ring;
        ring_ready_ = false;
        if (usable_ && arena_ptr_) {
            cudaSetDevice(device_);
            teardown_multicast();
        }
        if (!usable_ || !started_ || ring.size() < 2 || ring.size() > PCCL_K_MAX_RANKS) return;
        cudaSetDevice(device_);
        if (!ensure_arena()) return;
        // forget peers that left
        for (auto it = peers_.begin(); it != peers_.end();) {
            if (std::find(ring.begin(), ring.end(), it->first) == ring.end()) {
                unmap_peer(it->second);
                it = peers_.erase(it);
            } else {
                ++it;
            }
        }
        bool all = true;
        for (const auto &u : ring) {
            if (u == self_ || peers_.count(u)) continue;
            Peer p;
            if (fetch_peer(u, p)) {
                peers_.emplace(u, p);
            } else {
                all = false;
                break; // a peer on another host / without a GPU: this ring uses the TCP plane
            }
        }
        if (!all) return;
        // same geometry everywhere, one GPU per peer (kernels of two ranks that wait for each
        // other must never share a GPU)
        std::vector<std::array<uint8_t, 16>> gpus;
        std::array<uint8_t, 16> mine;
        std::memcpy(mine.data(), gpu_uuid_, 16);
        gpus.push_back(mine);
        for (const auto &[u, p] : peers_) {
            if (p.bytes != arena_bytes_ || p.slot_bytes != slot_bytes_ || p.n_slots != n_slots_) return;
            std::array<uint8_t, 16> g;
            std::memcpy(g.data(), p.gpu_uuid, 16);
            if (std::find(gpus.begin(), gpus.end(), g) != gpus.end()) {
                PLOG(INFO) << "two peers of the ring share one GPU: device plane disabled for this ring";
                return;
            }
            gpus.push_back(g);
        }
        ring_ready_ = true;
        PLOG(INFO) << "device plane ready: " << ring.size() << " peers mapped over NVLink";
        setup_multicast(ring);
    }

    uint32_t capabilities(const DeviceReduceArgs &a) override {
        std::lock_guard<std::mutex> lock(mu_);
        if (!ring_ready_ || a.world != ring_.size()) return 0;
        const PointerInfo s = classify_pointer(a.src), d = classify_pointer(a.dst);
        if (!s.is_device || !d.is_device || s.device != device_ || d.device != device_) return 0;
        if (a.dtype == wire::DType::F8E4M3) return 0;
        if (a.outer) {
            // fused outer step: fp32 average, all three tensors on this GPU, and the pseudo-gradient has to fit
            // into one staging slot (with several launches an abort could leave a half-applied update behind)
            if (a.dtype != wire::DType::F32 || a.op != wire::RedOp::Avg || a.quant != wire::QuantAlgo::None) return 0;
            for (const void *q : {static_cast<const void *>(a.outer->params), static_cast<const void *>(a.outer->inner),
                                  static_cast<const void *>(a.outer->momentum)}) {
                const PointerInfo i = classify_pointer(q);
                if (!i.is_device || i.device != device_) return 0;
            }
            if (pcclKernelP2PStagingBytes(a.count, pcclFloat, 0) > slot_bytes_) return 0;
        }
        switch (a.quant) {
        case wire::QuantAlgo::None:
            // Through the switch every GPU moves about N(1+1/P) bytes per direction (its own stores
            // are echoed back), peer-to-peer two-shot moves 2N(P-1)/P: multicast pays off from 4 peers.
            return kCapP2P | (mc_ready_ && a.world >= nvls_min_world_ && a.count * wire::dtype_size(a.dtype) >= nvls_min_bytes_ &&
                                      pcclKernelNvlsSupported(pccl_dtype(a.dtype), pccl_op(a.op))
                                  ? kCapNvls
                                  : 0u);
        // on the device plane uint8 min-max is block-scaled (128 elements per (min,max) pair)
        case wire::QuantAlgo::MinMax:
        case wire::QuantAlgo::BlockMinMax: return a.dtype == wire::DType::F32 && a.quant_dtype == wire::DType::U8 ? kCapP2P : 0u;
        case wire::QuantAlgo::BlockFp8: return a.dtype == wire::DType::F32 && a.quant_dtype == wire::DType::F8E4M3 ? kCapP2P : 0u;
        default: return 0;
        }
    }

    void announce(uint64_t seq) override {
        std::lock_guard<std::mutex> lock(order_mu_);
        order_.push_back(seq);
    }

    void cancel(uint64_t seq) override {
        {
            std::lock_guard<std::mutex> lock(order_mu_);
            order_.erase(std::remove(order_.begin(), order_.end(), seq), order_.end());
        }
        order_cv_.notify_all();
    }

    DeviceReduceResult reduce(const DeviceReduceArgs &a, const std::function<bool()> &abort_requested) override {
        DeviceReduceResult res;
        // Wait for this operation's turn (see announce()); leave the queue on every exit.  Device
        // operations run strictly one after the other, in the same order on every peer: two
        // collective kernels that each want (nearly) every SM and each wait for their twin on the
        // other GPUs would deadlock if two GPUs started them in different orders.
        struct Turn {
            CudaPlane *p;
            uint64_t seq;
            ~Turn() { p->cancel(seq); }
        } turn{this, a.seq};
        {
            std::unique_lock<std::mutex> lock(order_mu_);
            auto &q = order_;
            order_cv_.wait(lock, [&] { return q.empty() || q.front() == a.seq || std::find(q.begin(), q.end(), a.seq) == q.end(); });
        }
        std::vector<void *> arenas;
        void *mc_view = nullptr;
        {
            std::lock_guard<std::mutex> lock(mu_);
            if (!ring_ready_ || a.world != ring_.size()) return res;
            if (a.use_nvls) {
                if (!mc_ready_) return res;
                mc_view = reinterpret_cast<void *>(mc_.ptr);
            }
            for (const auto &u : ring_) arenas.push_back(u == self_ ? reinterpret_cast<void *>(arena_ptr_) : reinterpret_cast<void *>(peers_.at(u).ptr));
        }
        cudaSetDevice(device_);
        Slot &slot = slots_[a.seq % n_slots_];
        std::lock_guard<std::mutex> slot_lock(slot.mu); // one operation per slot at a time
        const size_t esz = wire::dtype_size(a.dtype);
        const int dtype = pccl_dtype(a.dtype), op = pccl_op(a.op), quant = kernel_quant(a.quant);
        const bool in_place = a.src == a.dst && !a.outer;

        // symmetric memory: the buffer itself is peer-visible, no staging copy
        const auto base = static_cast<uintptr_t>(arena_ptr_);
        const auto p = reinterpret_cast<uintptr_t>(a.src);
        const bool symmetric = in_place && quant == 0 && sym_bytes_ > 0 && p >= base + sym_offset() && p + a.count * esz <= base + arena_bytes_;

        // rounds: the staging slot bounds how much one launch can move
        uint64_t per_round = a.count;
        if (!symmetric) {
            const size_t need = pcclKernelP2PStagingBytes(a.count, dtype, quant);
            if (need > slot_bytes_) {
                const uint64_t unit = 32 * 4 * a.world * 16; // keeps chunk and quantisation-block alignment
                per_round = (slot_bytes_ - 4096) / (quant ? 2 : esz);
                per_round = std::max<uint64_t>(unit, per_round / unit * unit);
            }
        }
        const uint64_t n_rounds = (a.count + per_round - 1) / per_round;
        if (n_rounds > 255 || (a.outer && n_rounds != 1)) return res;

        if (in_place && !symmetric) {
            // keep the input for a rollback (the staging slot only holds the last round)
            if (slot.backup_bytes < a.count * esz) {
                if (slot.backup) cudaFree(slot.backup);
                slot.backup = nullptr;
                slot.backup_bytes = 0;
                if (cudaMalloc(&slot.backup, a.count * esz) != cudaSuccess) {
                    (void) cudaGetLastError();
                    return res;
                }
                slot.backup_bytes = a.count * esz;
            }
        }

        // inputs may still be in flight on the legacy default stream
        cudaEventRecord(slot.ready, cudaStreamLegacy);
        cudaStreamWaitEvent(slot.stream, slot.ready, 0);
        slot.host_words[0] = 0; // abort flag
        slot.host_words[1] = 0; // status
        cudaEventRecord(slot.start, slot.stream);
        if (in_place && !symmetric) cudaMemcpyAsync(slot.backup, a.src, a.count * esz, cudaMemcpyDeviceToDevice, slot.stream);
        slot.backup_valid = in_place && !symmetric;

        int algorithm = 0;
        // (opt-in, PCCL_PIPELINE=1: measured on 4xB200 it is no faster than the plain staged kernel --
        // 1.69-1.97 ms against 1.74 ms for 600 MB -- because the device-to-device copies compete with
        // the collective kernel for SMs; kept for machines where the copy engines take them)
        // Staged NVLS for ordinary device buffers, software-pipelined over segments: while segment s is
        // reduced through the switch (NVLink-bound, in place in the arena), segment s+1 is copied into
        // the arena and segment s-1 is copied out to the user buffer by the copy engines (HBM-bound).
        uint64_t seg_bytes = tune_ ? (std::max<uint64_t>(1, env_u64("PCCL_PIPELINE_SEGMENT_MB", 64)) << 20) : pipeline_seg_bytes_;
        seg_bytes = std::min<uint64_t>(seg_bytes, slot_bytes_ / 2) / 4096 * 4096;
        const bool pipelined = mc_view && !a.outer && !symmetric && quant == 0 && a.count * esz >= pipeline_min_bytes_ && seg_bytes >= (1u << 20) &&
                               (a.count * esz + seg_bytes - 1) / seg_bytes <= 250 && env_flag("PCCL_PIPELINE", false);
        if (pipelined) {
            const uint64_t bytes = a.count * esz;
            const uint64_t n_seg = (bytes + seg_bytes - 1) / seg_bytes, ring = slot_bytes_ / seg_bytes;
            for (auto *v : {&slot.ev_in, &slot.ev_kernel, &slot.ev_out})
                while (v->size() < n_seg) {
                    cudaEvent_t e;
                    if (cudaEventCreateWithFlags(&e, cudaEventDisableTiming) != cudaSuccess) return res;
                    v->push_back(e);
                }
            cudaStreamWaitEvent(slot.copy_in, slot.ready, 0);
            // the copy-in of this operation may not overtake the copy-out of the previous one
            cudaEventRecord(slot.ready, slot.stream);
            cudaStreamWaitEvent(slot.copy_in, slot.ready, 0);
            const auto *src8 = static_cast<const uint8_t *>(a.src);
            auto *dst8 = static_cast<uint8_t *>(a.dst);
            for (uint64_t sgm = 0; sgm < n_seg; ++sgm) {
                const uint64_t off = sgm * seg_bytes, n = std::min(seg_bytes, bytes - off);
                const uint64_t stage_off = slot_offset(a.seq % n_slots_) + (sgm % ring) * seg_bytes;
                uint8_t *stage = reinterpret_cast<uint8_t *>(arena_ptr_) + stage_off;
                if (sgm >= ring) cudaStreamWaitEvent(slot.copy_in, slot.ev_out[sgm - ring], 0);
                cudaMemcpyAsync(stage, src8 + off, n, cudaMemcpyDeviceToDevice, slot.copy_in);
                cudaEventRecord(slot.ev_in[sgm], slot.copy_in);
                cudaStreamWaitEvent(slot.stream, slot.ev_in[sgm], 0);
                pcclKernelNvlsArgs_t nk{};
                pcclKernelP2PArgs_t &k = nk.p2p;
                for (size_t i = 0; i < arenas.size(); ++i) k.arena[i] = arenas[i];
                k.rank = a.rank;
                k.world = a.world;
                k.flag_slot = static_cast<uint32_t>(a.seq % n_slots_);
                k.flag_base = (a.seq + 1) * 1024 + sgm * 4;
                k.data_offset = stage_off;
                k.data_capacity = seg_bytes + 512;
                k.src = k.dst = stage;
                k.count = n / esz;
                k.dtype = dtype;
                k.op = op;
                k.max_blocks = tune_ ? static_cast<int>(env_u64("PCCL_DEVICE_MAX_BLOCKS", 0)) : max_blocks_;
                k.src_in_arena = 1;
                k.abort_flag = reinterpret_cast<volatile uint32_t *>(slot.host_words);
                k.status = slot.host_words + 1;
                k.timeout_cycles = timeout_ms_ * static_cast<uint64_t>(clock_khz_);
                nk.mc_arena = mc_view;
                const int rc = pcclKernelAllReduceNvls(&nk, slot.stream);
                if (rc != 0) {
                    PLOG(ERR) << "all-reduce kernel launch failed: " << cudaGetErrorString(static_cast<cudaError_t>(rc));
                    cudaStreamSynchronize(slot.stream);
                    return res;
                }
                cudaEventRecord(slot.ev_kernel[sgm], slot.stream);
                cudaStreamWaitEvent(slot.copy_out, slot.ev_kernel[sgm], 0);
                cudaMemcpyAsync(dst8 + off, stage, n, cudaMemcpyDeviceToDevice, slot.copy_out);
                cudaEventRecord(slot.ev_out[sgm], slot.copy_out);
            }
            cudaStreamWaitEvent(slot.stream, slot.ev_out[n_seg - 1], 0);
            algorithm = 3;
            res.kernel_launches = static_cast<uint32_t>(n_seg);
        }
        for (uint64_t r = 0; r < (pipelined ? 0 : n_rounds); ++r) {
            const uint64_t off = r * per_round, n = std::min(per_round, a.count - off);
            pcclKernelP2PArgs_t k{};
            for (size_t i = 0; i < arenas.size(); ++i) k.arena[i] = arenas[i];
            k.rank = a.rank;
            k.world = a.world;
            k.flag_slot = static_cast<uint32_t>(a.seq % n_slots_);
            k.flag_base = (a.seq + 1) * 1024 + r * 4;
            k.data_offset = symmetric ? (p - base) + off * esz : slot_offset(a.seq % n_slots_);
            k.data_capacity = symmetric ? n * esz + 512 : slot_bytes_;
            k.src = static_cast<const uint8_t *>(a.src) + off * esz;
            k.dst = static_cast<uint8_t *>(a.dst) + off * esz;
            k.count = n;
            k.dtype = dtype;
            k.op = op;
            k.quant_algorithm = quant;
            k.algorithm = 0;
            k.max_blocks = max_blocks_;
            if (tune_) { // re-read the tuning knobs on every call (benchmark sweeps)
                k.max_blocks = static_cast<int>(env_u64("PCCL_DEVICE_MAX_BLOCKS", 0));
                k.algorithm = static_cast<int>(env_u64("PCCL_DEVICE_VARIANT", 0));
            }
            k.src_in_arena = symmetric ? 1 : 0;
            k.abort_flag = reinterpret_cast<volatile uint32_t *>(slot.host_words);
            k.status = slot.host_words + 1;
            k.timeout_cycles = timeout_ms_ * static_cast<uint64_t>(clock_khz_);
            int rc;
            // warp-specialised pipeline (copy-in || switch reduction || copy-out) for staged operands of one round
            const bool kernel_pipeline = mc_view && !symmetric && quant == 0 && n_rounds == 1 && n * esz >= pipeline_min_bytes_ &&
                                         (tune_ ? env_flag("PCCL_KERNEL_PIPELINE", kernel_pipeline_) : kernel_pipeline_);
            const int segments = static_cast<int>(tune_ ? env_u64("PCCL_PIPELINE_SEGMENTS", 0) : pipeline_segments_);
            if (a.outer) {
                pcclKernelOuterStep_t step{};
                step.inner_in = step.inner_out = a.outer->inner;
                step.params = a.outer->params;
                step.momentum = a.outer->momentum;
                step.lr = a.outer->lr;
                step.mu = a.outer->mu;
                step.first = a.outer->first ? 1 : 0;
                if (mc_view) {
                    pcclKernelNvlsArgs_t nk{};
                    nk.p2p = k;
                    nk.mc_arena = mc_view;
                    rc = kernel_pipeline ? pcclKernelAllReduceNvlsPipelined(&nk, &step, segments, slot.stream)
                                         : pcclKernelOuterStepNvls(&nk, &step, slot.stream);
                    algorithm = 3;
                } else {
                    rc = pcclKernelOuterStepP2P(&k, &step, slot.stream);
                    algorithm = 2;
                }
            } else if (mc_view) {
                pcclKernelNvlsArgs_t nk{};
                nk.p2p = k;
                nk.mc_arena = mc_view;
                rc = kernel_pipeline ? pcclKernelAllReduceNvlsPipelined(&nk, nullptr, segments, slot.stream)
                                     : pcclKernelAllReduceNvls(&nk, slot.stream);
                algorithm = 3;
            } else {
                rc = pcclKernelAllReduceP2P(&k, &algorithm, slot.stream);
            }
            if (rc != 0) {
                PLOG(ERR) << "all-reduce kernel launch failed: " << cudaGetErrorString(static_cast<cudaError_t>(rc));
                cudaStreamSynchronize(slot.stream);
                return res;
            }
        }
        cudaEventRecord(slot.stop, slot.stream);
        if (!pipelined) res.kernel_launches = static_cast<uint32_t>(n_rounds);

        // wait, relaying aborts from the control plane into the kernel
        int polls = 0;
        for (;;) {
            const cudaError_t q = cudaEventQuery(slot.stop);
            if (q == cudaSuccess) break;
            if (q != cudaErrorNotReady) {
                PLOG(ERR) << "device all-reduce failed: " << cudaGetErrorString(q);
                (void) cudaGetLastError();
                return res;
            }
            if ((++polls & 63) == 0 && abort_requested && abort_requested()) {
                reinterpret_cast<volatile uint32_t *>(slot.host_words)[0] = 1;
                res.aborted = true;
            }
            if (polls > 2000) std::this_thread::sleep_for(std::chrono::microseconds(50));
        }
        const uint32_t status = reinterpret_cast<volatile uint32_t *>(slot.host_words)[1];
        cudaEventElapsedTime(&res.device_ms, slot.start, slot.stop);
        res.plane = algorithm == 1 ? pcclExtPlaneP2POneShot : (algorithm == 3 ? pcclExtPlaneNvls : pcclExtPlaneP2PTwoShot);
        const uint64_t wire_bytes = quant ? a.count : a.count * esz;
        res.tx_bytes = res.rx_bytes = algorithm == 1   ? wire_bytes * (a.world - 1)
                                      : algorithm == 3 ? wire_bytes + wire_bytes / a.world
                                                       : 2 * wire_bytes * (a.world - 1) / a.world;
        res.ok = status == 0;
        if (status == 2) PLOG(ERR) << "device all-reduce timed out waiting for a peer";
        return res;
    }

    bool restore_input(const DeviceReduceArgs &a) override {
        if (!usable_ || slots_.empty()) return false;
        cudaSetDevice(device_);
        Slot &slot = slots_[a.seq % n_slots_];
        std::lock_guard<std::mutex> slot_lock(slot.mu);
        if (!slot.backup_valid || !slot.backup) return false;
        const size_t bytes = a.count * wire::dtype_size(a.dtype);
        if (cudaMemcpyAsync(a.dst, slot.backup, bytes, cudaMemcpyDeviceToDevice, slot.stream) != cudaSuccess) return false;
        return cudaStreamSynchronize(slot.stream) == cudaSuccess;
    }

    bool copy_to_host(void *host_dst, const void *dev_src, size_t bytes) override { return copy(host_dst, dev_src, bytes, cudaMemcpyDeviceToHost); }
    bool copy_to_device(void *dev_dst, const void *host_src, size_t bytes) override { return copy(dev_dst, host_src, bytes, cudaMemcpyHostToDevice); }

    bool hash_device(const void *dev_ptr, size_t bytes, wire::HashType type, uint64_t &out) override {
        if (!cuda_runtime_usable()) return false;
        const PointerInfo info = classify_pointer(dev_ptr);
        int prev = -1;
        cudaGetDevice(&prev);
        if (info.is_device && info.device >= 0) cudaSetDevice(info.device);
        const int rc = pcclKernelHashSync(dev_ptr, bytes, type == wire::HashType::Simple ? 0 : 1, &out);
        if (prev >= 0) cudaSetDevice(prev);
        if (rc != 0) PLOG(ERR) << "device hash failed: " << cudaGetErrorString(static_cast<cudaError_t>(rc));
        return rc == 0;
    }

    bool pseudo_grad(float *out, const float *params, const float *inner, size_t count) override {
        return elementwise([&](cudaStream_t st) { return pcclKernelPseudoGrad(out, params, inner, count, st); });
    }
    bool outer_nesterov(const OuterStepArgs &step, const float *grad, size_t count) override {
        return elementwise([&](cudaStream_t st) {
            return pcclKernelOuterNesterov(step.params, step.momentum, grad, step.inner, count, step.lr, step.mu, step.first ? 1 : 0, st);
        });
    }

    void *fuse_buffer(size_t bytes) override {
        if (!usable_) return nullptr;
        cudaSetDevice(device_);
        std::lock_guard<std::mutex> lock(fuse_mu_);
        // the fixed scratch region of the symmetric heap when the operand fits: reduced in place, no staging
        if (bytes <= internal_bytes_) {
            std::lock_guard<std::mutex> arena_lock(mu_);
            if (ensure_arena()) return reinterpret_cast<void *>(arena_ptr_ + sym_offset() + (sym_bytes_ - internal_bytes_));
        }
        if (fuse_ptr_ && fuse_bytes_ >= bytes) return fuse_ptr_;
        if (fuse_ptr_) cudaFree(fuse_ptr_);
        fuse_ptr_ = nullptr;
        fuse_bytes_ = 0;
        // grow generously: parameter lists do not change from step to step
        const size_t want = round_up(bytes + bytes / 8, 2u << 20);
        if (cudaMalloc(&fuse_ptr_, want) != cudaSuccess) {
            (void) cudaGetLastError();
            fuse_ptr_ = nullptr;
            return nullptr;
        }
        fuse_bytes_ = want;
        return fuse_ptr_;
    }
    std::unique_lock<std::mutex> fuse_lock() override { return std::unique_lock<std::mutex>(fuse_use_mu_); }

    bool gather(const std::vector<Part> &parts, void *fused, bool into_fused) override {
        if (!usable_) return false;
        cudaSetDevice(device_);
        std::lock_guard<std::mutex> lock(copy_mu_);
        if (!util_stream_) RT_TRY(cudaStreamCreateWithFlags(&util_stream_, cudaStreamNonBlocking));
        cudaEvent_t ready;
        RT_TRY(cudaEventCreateWithFlags(&ready, cudaEventDisableTiming));
        cudaEventRecord(ready, cudaStreamLegacy);
        cudaStreamWaitEvent(util_stream_, ready, 0);
        cudaEventDestroy(ready);
        auto *base = static_cast<uint8_t *>(fused);
        size_t off = 0;
        for (const Part &p : parts) {
            if (into_fused)
                RT_TRY(cudaMemcpyAsync(base + off, p.ptr, p.bytes, cudaMemcpyDeviceToDevice, util_stream_));
            else
                RT_TRY(cudaMemcpyAsync(p.ptr, base + off, p.bytes, cudaMemcpyDeviceToDevice, util_stream_));
            off += round_up(p.bytes, 16);
        }
        RT_TRY(cudaStreamSynchronize(util_stream_));
        return true;
    }

    bool peer_mapped(const Uuid &peer) override {
        std::lock_guard<std::mutex> lock(mu_);
        return ring_ready_ && peers_.count(peer) > 0;
    }
    size_t exchange_bytes() override {
        std::lock_guard<std::mutex> lock(mu_);
        return arena_ptr_ && !slots_.empty() ? slot_bytes_ : 0;
    }
    std::unique_lock<std::mutex> exchange_lock() override { return std::unique_lock<std::mutex>(slots_[0].mu); }
    bool export_chunk(const void *src, bool, size_t bytes) override {
        if (bytes > slot_bytes_) return false;
        return copy(reinterpret_cast<void *>(arena_ptr_ + slot_offset(0)), src, bytes, cudaMemcpyDefault);
    }
    bool import_chunk(const Uuid &from, void *dst, bool, size_t bytes) override {
        CUdeviceptr base = 0;
        {
            std::lock_guard<std::mutex> lock(mu_);
            const auto it = peers_.find(from);
            if (it == peers_.end() || bytes > slot_bytes_) return false;
            base = it->second.ptr;
        }
        return copy(dst, reinterpret_cast<const void *>(base + slot_offset(0)), bytes, cudaMemcpyDefault);
    }

    void *host_alloc(size_t bytes) override {
        if (cuda_runtime_usable()) {
            void *p = nullptr;
            if (cudaHostAlloc(&p, bytes, cudaHostAllocPortable) == cudaSuccess) return p;
            (void) cudaGetLastError();
            return nullptr;
        }
        return std::malloc(bytes);
    }
    void host_free(void *ptr) override {
        if (!ptr) return;
        if (cuda_runtime_usable())
            cudaFreeHost(ptr);
        else
            std::free(ptr);
    }

    void *symmetric_alloc(size_t bytes) override {
        std::lock_guard<std::mutex> lock(mu_);
        if (!usable_ || sym_bytes_ == 0) return nullptr;
        cudaSetDevice(device_);
        if (!ensure_arena()) return nullptr;
        bytes = round_up(bytes, 512);
        // first fit over the gaps between live allocations (same call order on every peer gives
        // the same offsets, which is what makes the memory "symmetric")
        size_t cursor = 0;
        for (const auto &[off, len] : sym_allocs_) {
            if (off - cursor >= bytes) break;
            cursor = off + len;
        }
        if (cursor + bytes > sym_bytes_ - internal_bytes_) return nullptr; // the tail belongs to fuse_buffer()
        sym_allocs_[cursor] = bytes;
        return reinterpret_cast<void *>(arena_ptr_ + sym_offset() + cursor);
    }

    bool symmetric_free(void *ptr) override {
        std::lock_guard<std::mutex> lock(mu_);
        if (!arena_ptr_) return false;
        const auto off = reinterpret_cast<uintptr_t>(ptr) - (arena_ptr_ + sym_offset());
        return sym_allocs_.erase(off) > 0;
    }

private:
    struct Peer {
        CUdeviceptr ptr = 0;
        size_t bytes = 0, slot_bytes = 0;
        uint32_t n_slots = 0;
        CUmemGenericAllocationHandle handle = 0;
        uint8_t gpu_uuid[16] = {0};
    };
    struct Slot {
        std::mutex mu;
        cudaStream_t stream = nullptr;
        cudaEvent_t start = nullptr, stop = nullptr, ready = nullptr;
        cudaStream_t copy_in = nullptr, copy_out = nullptr; // staged NVLS pipeline
        std::vector<cudaEvent_t> ev_in, ev_kernel, ev_out;
        uint32_t *host_words = nullptr; // [0] abort flag, [1] status; host-mapped
        void *backup = nullptr;
        size_t backup_bytes = 0;
        bool backup_valid = false;
    };

    size_t slot_offset(uint64_t slot) const { return pcclKernelP2PControlBytes() + slot * slot_bytes_; }
    size_t sym_offset() const { return pcclKernelP2PControlBytes() + size_t(n_slots_) * slot_bytes_; }

    static int pccl_dtype(wire::DType t) {
        switch (t) {
        case wire::DType::U8: return pcclUint8;
        case wire::DType::I8: return pcclInt8;
        case wire::DType::U16: return pcclUint16;
        case wire::DType::I16: return pcclInt16;
        case wire::DType::U32: return pcclUint32;
        case wire::DType::I32: return pcclInt32;
        case wire::DType::U64: return pcclUint64;
        case wire::DType::I64: return pcclInt64;
        case wire::DType::F16: return pcclFloat16;
        case wire::DType::BF16: return pcclBFloat16;
        case wire::DType::F32: return pcclFloat;
        case wire::DType::F64: return pcclDouble;
        default: return -1;
        }
    }
    static int pccl_op(wire::RedOp op) {
        switch (op) {
        case wire::RedOp::Sum: return pcclSum;
        case wire::RedOp::Avg: return pcclAvg;
        case wire::RedOp::Prod: return pcclProd;
        case wire::RedOp::Max: return pcclMax;
        case wire::RedOp::Min: return pcclMin;
        default: return -1;
        }
    }
    static int kernel_quant(wire::QuantAlgo q) {
        switch (q) {
        case wire::QuantAlgo::MinMax:
        case wire::QuantAlgo::BlockMinMax: return pcclQuantBlockMinMax;
        case wire::QuantAlgo::BlockFp8: return pcclQuantBlockFp8;
        default: return pcclQuantNone;
        }
    }

    // Device current.  Has its own lock: the rendezvous server thread must be able to create and
    // export the arena while the main thread sits in on_ring_changed() waiting for a PEER's server
    // (two peers fetching each other's arenas would otherwise deadlock until the timeout).
    bool ensure_arena() {
        std::lock_guard<std::mutex> arena_lock(arena_mu_);
        if (arena_ptr_) return true;
        const Driver &d = driver();
        CUmemAllocationProp prop{};
        prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
        prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
        prop.location.id = device_;
        prop.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
        size_t gran = 2u << 20;
        CU_TRY(d.MemGetAllocationGranularity(&gran, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED));
        if (d.multicast) {
            CUmulticastObjectProp mprop{};
            mprop.numDevices = 2;
            mprop.size = gran;
            mprop.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
            size_t mg = 0;
            if (d.MulticastGetGranularity(&mg, &mprop, CU_MULTICAST_GRANULARITY_RECOMMENDED) == CUDA_SUCCESS && mg > 0) {
                mc_gran_ = mg;
                gran = std::max(gran, mg);
            }
        }
        arena_bytes_ = round_up(pcclKernelP2PControlBytes() + size_t(n_slots_) * slot_bytes_ + sym_bytes_, gran);
        CU_TRY(d.MemCreate(&arena_handle_, arena_bytes_, &prop, 0));
        CU_TRY(d.MemAddressReserve(&arena_ptr_, arena_bytes_, gran, 0, 0));
        CU_TRY(d.MemMap(arena_ptr_, arena_bytes_, 0, arena_handle_, 0));
        CUmemAccessDesc acc{};
        acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
        acc.location.id = device_;
        acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
        CU_TRY(d.MemSetAccess(arena_ptr_, arena_bytes_, &acc, 1));
        RT_TRY(cudaMemset(reinterpret_cast<void *>(arena_ptr_), 0, pcclKernelP2PControlBytes()));
        RT_TRY(cudaDeviceSynchronize());
        slots_ = std::vector<Slot>(n_slots_);
        for (auto &s : slots_) {
            RT_TRY(cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking));
            RT_TRY(cudaStreamCreateWithFlags(&s.copy_in, cudaStreamNonBlocking));
            RT_TRY(cudaStreamCreateWithFlags(&s.copy_out, cudaStreamNonBlocking));
            RT_TRY(cudaEventCreate(&s.start));
            RT_TRY(cudaEventCreate(&s.stop));
            RT_TRY(cudaEventCreateWithFlags(&s.ready, cudaEventDisableTiming));
            RT_TRY(cudaHostAlloc(reinterpret_cast<void **>(&s.host_words), 64, cudaHostAllocMapped));
            std::memset(s.host_words, 0, 64);
        }
        PLOG(INFO) << "device arena of " << (arena_bytes_ >> 20) << " MiB on GPU " << device_;
        return true;
    }

    void unmap_peer(Peer &p) {
        if (!p.ptr) return;
        driver().MemUnmap(p.ptr, p.bytes);
        driver().MemAddressFree(p.ptr, p.bytes);
        driver().MemRelease(p.handle);
        p.ptr = 0;
    }

    // mu_ held, device current
    bool fetch_peer(const Uuid &who, Peer &out) {
        const int sock = ::socket(AF_UNIX, SOCK_STREAM | SOCK_CLOEXEC, 0);
        if (sock < 0) return false;
        socklen_t len;
        const sockaddr_un sa = rendezvous_address(who, len);
        bool connected = false;
        for (int attempt = 0; attempt < 40 && !connected; ++attempt) {
            if (::connect(sock, reinterpret_cast<const sockaddr *>(&sa), len) == 0)
                connected = true;
            else if (attempt < 2 || errno != ECONNREFUSED)
                std::this_thread::sleep_for(std::chrono::milliseconds(25));
            else
                break; // nobody listens under that name: the peer is not on this host
        }
        if (!connected) {
            ::close(sock);
            return false;
        }
        RendezvousRequest req{kMagic, kReqArena, {0}, 0};
        std::memcpy(req.requester, self_.data.data(), 16);
        RendezvousReply rep{};
        int fd = -1;
        const bool ok = ::send(sock, &req, sizeof req, MSG_NOSIGNAL) == static_cast<ssize_t>(sizeof req) &&
                        recv_with_fd(sock, &rep, sizeof rep, fd, 20000) && rep.magic == kMagic && rep.status == 0 && fd >= 0;
        ::close(sock);
        if (!ok) {
            if (fd >= 0) ::close(fd);
            return false;
        }
        const Driver &d = driver();
        Peer p;
        p.bytes = rep.arena_bytes;
        p.slot_bytes = rep.slot_bytes;
        p.n_slots = rep.n_slots;
        std::memcpy(p.gpu_uuid, rep.gpu_uuid, 16);
        const CUresult imp = d.MemImportFromShareableHandle(&p.handle, reinterpret_cast<void *>(static_cast<uintptr_t>(fd)), CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
        ::close(fd);
        if (imp != CUDA_SUCCESS) {
            PLOG(WARN) << "cannot import the arena of peer " << who.str() << " (CUresult " << static_cast<int>(imp) << ")";
            return false;
        }
        if (d.MemAddressReserve(&p.ptr, p.bytes, 2u << 20, 0, 0) != CUDA_SUCCESS) {
            d.MemRelease(p.handle);
            return false;
        }
        CUmemAccessDesc acc{};
        acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
        acc.location.id = device_;
        acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
        if (d.MemMap(p.ptr, p.bytes, 0, p.handle, 0) != CUDA_SUCCESS || d.MemSetAccess(p.ptr, p.bytes, &acc, 1) != CUDA_SUCCESS) {
            PLOG(WARN) << "cannot map the arena of peer " << who.str() << " (no peer access between the GPUs?)";
            d.MemUnmap(p.ptr, p.bytes);
            d.MemAddressFree(p.ptr, p.bytes);
            d.MemRelease(p.handle);
            return false;
        }
        out = p;
        return true;
    }

    static uint64_t ring_key_of(const std::vector<Uuid> &ring) {
        uint64_t h = 1469598103934665603ull;
        for (const auto &u : ring)
            for (uint8_t b : u.data) h = (h ^ b) * 1099511628211ull;
        return h ? h : 1;
    }

    // One request/reply with a peer's rendezvous server.  Returns false when unreachable.
    bool ask(const Uuid &who, uint32_t kind, uint64_t ring_key, RendezvousReply &rep, int &fd) {
        fd = -1;
        const int sock = ::socket(AF_UNIX, SOCK_STREAM | SOCK_CLOEXEC, 0);
        if (sock < 0) return false;
        socklen_t len;
        const sockaddr_un sa = rendezvous_address(who, len);
        if (::connect(sock, reinterpret_cast<const sockaddr *>(&sa), len) != 0) {
            ::close(sock);
            return false;
        }
        RendezvousRequest req{kMagic, kind, {0}, ring_key};
        std::memcpy(req.requester, self_.data.data(), 16);
        const bool ok = ::send(sock, &req, sizeof req, MSG_NOSIGNAL) == static_cast<ssize_t>(sizeof req) &&
                        recv_with_fd(sock, &rep, sizeof rep, fd, 10000) && rep.magic == kMagic;
        ::close(sock);
        return ok;
    }

    // mu_ held, device current
    void teardown_multicast() {
        mc_ready_ = false;
        std::lock_guard<std::mutex> lock(mc_mu_);
        if (mc_.ptr) {
            driver().MemUnmap(mc_.ptr, arena_bytes_);
            driver().MemAddressFree(mc_.ptr, arena_bytes_);
        }
        if (mc_.bound) {
            CUdevice dev;
            if (driver().DeviceGet(&dev, device_) == CUDA_SUCCESS) driver().MulticastUnbind(mc_.handle, dev, 0, arena_bytes_);
        }
        if (mc_.have) driver().MemRelease(mc_.handle);
        mc_ = Multicast{};
    }

    // Builds the multicast view of all arenas of the ring (NVSwitch multimem path).  Every member
    // runs this after the unicast mappings exist: the first member of the ring creates the
    // object, everybody adds its device, and only when all devices are in does anybody bind
    // memory (a CUDA requirement).  Failure just leaves the NVLS path off; whether it is used is
    // agreed per operation through the master, so a member without it cannot diverge.
    // mu_ held, device current
    void setup_multicast(const std::vector<Uuid> &ring) {
        const Driver &d = driver();
        if (!d.multicast || !env_flag("PCCL_NVLS", true)) return;
        CUdevice dev;
        int supported = 0;
        if (d.DeviceGet(&dev, device_) != CUDA_SUCCESS ||
            d.DeviceGetAttribute(&supported, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev) != CUDA_SUCCESS || !supported) {
            PLOG(INFO) << "multicast is not supported on this device: NVLS path off";
            return;
        }
        if (mc_gran_ == 0 || arena_bytes_ % mc_gran_ != 0) return;
        const uint64_t key = ring_key_of(ring);
        const Uuid &leader = ring[0];
        const uint32_t world = static_cast<uint32_t>(ring.size());
        CUmulticastObjectProp prop{};
        prop.numDevices = world;
        prop.size = arena_bytes_;
        prop.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
        const auto deadline = clock_type::now() + std::chrono::seconds(15);
        CUmemGenericAllocationHandle handle = 0;
        if (leader == self_) {
            const CUresult rc = d.MulticastCreate(&handle, &prop);
            if (rc != CUDA_SUCCESS) {
                PLOG(INFO) << "cuMulticastCreate failed (" << static_cast<int>(rc) << "): NVLS path off";
                return;
            }
            std::lock_guard<std::mutex> lock(mc_mu_);
            mc_.ring_key = key;
            mc_.handle = handle;
            mc_.created = mc_.have = true;
        } else {
            bool got = false;
            while (!got && clock_type::now() < deadline) {
                RendezvousReply rep{};
                int fd = -1;
                if (ask(leader, kReqMulticast, key, rep, fd) && rep.status == 0 && fd >= 0) {
                    const CUresult rc = d.MemImportFromShareableHandle(&handle, reinterpret_cast<void *>(static_cast<uintptr_t>(fd)),
                                                                       CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
                    ::close(fd);
                    if (rc != CUDA_SUCCESS) return;
                    got = true;
                } else {
                    if (fd >= 0) ::close(fd);
                    if (rep.status == 1) return; // leader cannot do multicast
                    std::this_thread::sleep_for(std::chrono::milliseconds(5));
                }
            }
            if (!got) return;
            std::lock_guard<std::mutex> lock(mc_mu_);
            mc_.ring_key = key;
            mc_.handle = handle;
            mc_.have = true;
        }
        if (d.MulticastAddDevice(handle, dev) != CUDA_SUCCESS) {
            PLOG(INFO) << "cuMulticastAddDevice failed: NVLS path off";
            return;
        }
        // tell the leader, then wait until every member has added its device
        if (leader == self_) {
            std::lock_guard<std::mutex> lock(mc_mu_);
            mc_.added++;
        } else {
            RendezvousReply rep{};
            int fd = -1;
            if (!ask(leader, kReqMulticastAdded, key, rep, fd) || rep.status != 0) return;
        }
        for (;;) {
            uint32_t added = 0;
            if (leader == self_) {
                std::lock_guard<std::mutex> lock(mc_mu_);
                added = mc_.added;
            } else {
                RendezvousReply rep{};
                int fd = -1;
                if (!ask(leader, kReqMulticastReady, key, rep, fd) || rep.status != 0) return;
                added = rep.added_count;
            }
            if (added >= world) break;
            if (clock_type::now() > deadline) {
                PLOG(WARN) << "multicast setup timed out (" << added << "/" << world << " devices): NVLS path off";
                return;
            }
            std::this_thread::sleep_for(std::chrono::milliseconds(2));
        }
        CUresult rc = d.MulticastBindMem(handle, 0, arena_handle_, 0, arena_bytes_, 0);
        if (rc != CUDA_SUCCESS) {
            PLOG(INFO) << "cuMulticastBindMem failed (" << static_cast<int>(rc) << "): NVLS path off";
            return;
        }
        CUdeviceptr ptr = 0;
        CUmemAccessDesc acc{};
        acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
        acc.location.id = device_;
        acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
        if (d.MemAddressReserve(&ptr, arena_bytes_, mc_gran_, 0, 0) != CUDA_SUCCESS) return;
        if (d.MemMap(ptr, arena_bytes_, 0, handle, 0) != CUDA_SUCCESS || d.MemSetAccess(ptr, arena_bytes_, &acc, 1) != CUDA_SUCCESS) {
            d.MemAddressFree(ptr, arena_bytes_);
            return;
        }
        {
            std::lock_guard<std::mutex> lock(mc_mu_);
            mc_.bound = true;
            mc_.ptr = ptr;
        }
        mc_ready_ = true;
        PLOG(INFO) << "NVSwitch multicast view of " << world << " arenas mapped: NVLS path on";
    }

    void serve() {
        cudaSetDevice(device_);
        while (!stopping_) {
            pollfd p{listen_fd_, POLLIN, 0};
            if (::poll(&p, 1, 100) <= 0) continue;
            const int c = ::accept4(listen_fd_, nullptr, nullptr, SOCK_CLOEXEC);
            if (c < 0) continue;
            RendezvousRequest req{};
            pollfd cp{c, POLLIN, 0};
            RendezvousReply rep{};
            rep.magic = kMagic;
            rep.status = 1;
            int fd = -1;
            const bool got = ::poll(&cp, 1, 2000) > 0 && ::recv(c, &req, sizeof req, MSG_WAITALL) == static_cast<ssize_t>(sizeof req) &&
                             req.magic == kMagic;
            if (got && req.kind == kReqArena) {
                if (ensure_arena() && driver().MemExportToShareableHandle(&fd, arena_handle_, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0) == CUDA_SUCCESS) {
                    rep.status = 0;
                    rep.arena_bytes = arena_bytes_;
                    rep.slot_bytes = slot_bytes_;
                    rep.sym_bytes = sym_bytes_;
                    rep.n_slots = n_slots_;
                    rep.pid = static_cast<uint32_t>(::getpid());
                    std::memcpy(rep.gpu_uuid, gpu_uuid_, 16);
                }
            } else if (got) {
                // multicast bookkeeping; this peer answers as the leader (first member of the ring)
                std::lock_guard<std::mutex> lock(mc_mu_);
                if (mc_.created && mc_.ring_key == req.ring_key) {
                    if (req.kind == kReqMulticast) {
                        if (driver().MemExportToShareableHandle(&fd, mc_.handle, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0) == CUDA_SUCCESS) rep.status = 0;
                    } else if (req.kind == kReqMulticastAdded) {
                        mc_.added++;
                        rep.status = 0;
                    } else if (req.kind == kReqMulticastReady) {
                        rep.status = 0;
                    }
                    rep.added_count = mc_.added;
                } else {
                    rep.status = 2; // not (yet) the leader of that epoch: ask again
                }
            }
            send_with_fd(c, &rep, sizeof rep, fd);
            if (fd >= 0) ::close(fd);
            ::close(c);
        }
    }

    // One element-wise launch on the utility stream, ordered after the legacy default stream, waited for.
    template <typename Launch> bool elementwise(Launch launch) {
        if (!usable_) return false;
        cudaSetDevice(device_);
        std::lock_guard<std::mutex> lock(copy_mu_);
        if (!util_stream_) RT_TRY(cudaStreamCreateWithFlags(&util_stream_, cudaStreamNonBlocking));
        cudaEvent_t ready;
        RT_TRY(cudaEventCreateWithFlags(&ready, cudaEventDisableTiming));
        cudaEventRecord(ready, cudaStreamLegacy);
        cudaStreamWaitEvent(util_stream_, ready, 0);
        cudaEventDestroy(ready);
        if (launch(util_stream_) != 0) return false;
        RT_TRY(cudaStreamSynchronize(util_stream_));
        return true;
    }

    bool copy(void *dst, const void *src, size_t bytes, cudaMemcpyKind kind) {
        if (!cuda_runtime_usable()) return false;
        if (kind == cudaMemcpyDefault) {
            cudaSetDevice(device_);
        } else {
            const PointerInfo info = classify_pointer(kind == cudaMemcpyDeviceToHost ? src : dst);
            if (info.is_device && info.device >= 0) cudaSetDevice(info.device);
        }
        std::lock_guard<std::mutex> lock(copy_mu_);
        if (!util_stream_) RT_TRY(cudaStreamCreateWithFlags(&util_stream_, cudaStreamNonBlocking));
        if (kind == cudaMemcpyDeviceToHost || kind == cudaMemcpyDefault) {
            // the source may still be written by work queued on the legacy default stream
            cudaEvent_t ready;
            RT_TRY(cudaEventCreateWithFlags(&ready, cudaEventDisableTiming));
            cudaEventRecord(ready, cudaStreamLegacy);
            cudaStreamWaitEvent(util_stream_, ready, 0);
            cudaEventDestroy(ready);
        }
        RT_TRY(cudaMemcpyAsync(dst, src, bytes, kind, util_stream_));
        RT_TRY(cudaStreamSynchronize(util_stream_));
        return true;
    }

    bool usable_ = false, started_ = false;
    std::atomic<bool> stopping_{false};
    int device_ = 0;
    int clock_khz_ = 1900000;
    uint8_t gpu_uuid_[16] = {0};
    Uuid self_;
    int listen_fd_ = -1;
    std::thread server_;

    std::mutex mu_;
    std::mutex arena_mu_; // guards arena creation only; never held while waiting for a peer
    struct Multicast {
        uint64_t ring_key = 0;
        CUmemGenericAllocationHandle handle = 0;
        bool created = false; // this peer created the object (leader)
        bool have = false, bound = false;
        uint32_t added = 0;
        CUdeviceptr ptr = 0;
    };
    std::mutex mc_mu_; // leader-side multicast bookkeeping, shared with the server thread
    Multicast mc_;
    bool mc_ready_ = false;
    size_t mc_gran_ = 0;
    size_t slot_bytes_ = 0, sym_bytes_ = 0, internal_bytes_ = 0, arena_bytes_ = 0;
    bool kernel_pipeline_ = false;
    uint64_t pipeline_segments_ = 0;
    uint32_t n_slots_ = 4;
    uint64_t timeout_ms_ = 30000;
    int max_blocks_ = 0;
    uint32_t nvls_min_world_ = 4;
    bool tune_ = false;
    uint64_t pipeline_min_bytes_ = 48u << 20, pipeline_seg_bytes_ = 64u << 20;
    uint64_t nvls_min_bytes_ = 1u << 20;
    CUdeviceptr arena_ptr_ = 0;
    CUmemGenericAllocationHandle arena_handle_ = 0;
    std::vector<Slot> slots_;
    std::map<size_t, size_t> sym_allocs_;
    std::unordered_map<Uuid, Peer> peers_;
    std::vector<Uuid> ring_;
    bool ring_ready_ = false;

    std::mutex order_mu_;
    std::condition_variable order_cv_;
    std::deque<uint64_t> order_; // announced device operations in Commence arrival order

    std::mutex copy_mu_;
    cudaStream_t util_stream_ = nullptr;

    std::mutex fuse_mu_, fuse_use_mu_;
    void *fuse_ptr_ = nullptr; // cudaMalloc'd fallback for operands beyond the scratch region
    size_t fuse_bytes_ = 0;
};

} // namespace

bool cuda_runtime_usable() {
    static const bool ok = [] {
        int n = 0;
        const cudaError_t rc = cudaGetDeviceCount(&n);
        if (rc != cudaSuccess) (void) cudaGetLastError();
        return rc == cudaSuccess && n > 0;
    }();
    return ok;
}

PointerInfo classify_pointer(const void *p) {
    PointerInfo info;
    if (!p || !cuda_runtime_usable()) return info;
    cudaPointerAttributes attr{};
    if (cudaPointerGetAttributes(&attr, p) != cudaSuccess) {
        (void) cudaGetLastError();
        return info;
    }
    if (attr.type == cudaMemoryTypeDevice || attr.type == cudaMemoryTypeManaged) {
        info.is_device = true;
        info.device = attr.device;
    }
    return info;
}

std::unique_ptr<Plane> Plane::create() { return std::make_unique<CudaPlane>(); }

} // namespace pccl::device
#endif // PCCL_NO_CUDA
