// CUDA implementation of the device plane (see plane.hpp for the design).
//
// Driver-API entry points (virtual memory management, multicast) are resolved at run time through
// cudaGetDriverEntryPoint, so libpccl.so has no link-time dependency on libcuda and loads on
// machines without a GPU.
#ifndef PCCL_NO_CUDA
#include <cuda.h>
#include <cuda_runtime.h>
#include <poll.h>
#include <sys/mman.h>
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
#include "sequencer.hpp"

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
enum : uint32_t { kReqArena = 1, kReqMulticast = 2, kReqMulticastAdded = 3, kReqMulticastReady = 4, kReqSequencer = 5 };

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
        slot_bytes_ = round_up(env_u64("PCCL_ARENA_SLOT_MB", 640) << 20, 2u << 20);
        n_slots_ = static_cast<uint32_t>(std::clamp<uint64_t>(env_u64("PCCL_ARENA_SLOTS", 2), 1, 15));
        // every flag bank owns a small double-buffered staging area for latency-bound operations
        small_half_ = round_up(std::clamp<uint64_t>(env_u64("PCCL_SMALL_STAGING_KB", 1024), 64, 16384) << 10, 4096);
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
        small_grid_ = static_cast<int>(std::clamp<uint64_t>(env_u64("PCCL_SMALL_OP_BLOCKS", 8), 1, 8));
        use_sequencer_ = env_flag("PCCL_SEQUENCER", true);
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
        for (auto &b : banks_) {
            if (b.stream) cudaStreamSynchronize(b.stream);
            if (b.stream) cudaStreamDestroy(b.stream);
            for (cudaEvent_t e : {b.start, b.stop, b.ready})
                if (e) cudaEventDestroy(e);
            if (b.host_words) cudaFreeHost(b.host_words);
            if (b.hash_host) cudaFreeHost(b.hash_host);
            if (b.hash_dev) cudaFree(b.hash_dev);
        }
        banks_.clear();
        for (auto &sl : slots_)
            if (sl.last_use) cudaEventDestroy(sl.last_use);
        slots_.clear();
        if (util_stream_) cudaStreamDestroy(util_stream_), util_stream_ = nullptr;
        if (arena_ptr_) {
            driver().MemUnmap(arena_ptr_, arena_bytes_);
            driver().MemAddressFree(arena_ptr_, arena_bytes_);
            driver().MemRelease(arena_handle_);
            arena_ptr_ = 0;
        }
        drop_sequencer();
    }

    void on_ring_changed(const std::vector<Uuid> &ring) override {
        std::lock_guard<std::mutex> lock(mu_);
        ring_ = ring;
        ring_ready_ = false;
        {
            std::lock_guard<std::mutex> l(launch_mu_);
            drop_sequencer();
            next_launch_ = 0;
            big_count_ = 0;
            last_big_bank_ = -1;
        }
        launch_cv_.notify_all();
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
        if (use_sequencer_) setup_sequencer(ring);
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
            return kCapP2P | (mc_ready_ && nvls_eligible_locked(a) ? kCapNvls : 0u);
        // on the device plane uint8 min-max is block-scaled (128 elements per (min,max) pair)
        case wire::QuantAlgo::MinMax:
        case wire::QuantAlgo::BlockMinMax: return a.dtype == wire::DType::F32 && a.quant_dtype == wire::DType::U8 ? kCapP2P : 0u;
        case wire::QuantAlgo::BlockFp8: return a.dtype == wire::DType::F32 && a.quant_dtype == wire::DType::F8E4M3 ? kCapP2P : 0u;
        // zero-point/scale with uint8 codes runs block-wise on the device; other code types take the TCP ring
        case wire::QuantAlgo::ZeroPointScale: return a.dtype == wire::DType::F32 && a.quant_dtype == wire::DType::U8 ? kCapP2P : 0u;
        default: return 0;
        }
    }

    bool lease_ready() override {
        std::lock_guard<std::mutex> lock(launch_mu_);
        return sequencer_.attached();
    }
    bool multicast_ready() override {
        std::lock_guard<std::mutex> lock(mu_);
        return mc_ready_;
    }
    bool nvls_eligible(const DeviceReduceArgs &a) override {
        std::lock_guard<std::mutex> lock(mu_);
        return nvls_eligible_locked(a);
    }

    void revoke_leased(uint64_t tag, bool all) override {
        std::lock_guard<std::mutex> lock(inflight_mu_);
        for (auto &[ticket, f] : inflight_) {
            if (!f.leased || (!all && f.tag != tag)) continue;
            f.revoked = true;
            reinterpret_cast<volatile uint32_t *>(banks_[f.bank].host_words)[0] = 1;
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

    // Launches the kernel(s) of one reduction and (unless enqueue_only) waits for them.
    //
    // Order.  Collective kernels of different operations must start in the same order on every peer (each wants SMs and
    // waits for its twins).  With the same-box sequencer attached, every operation -- voted at the master or leased --
    // takes its place from the shared log and launches strictly in that order; without it, voted operations run one
    // at a time in the order their Commence packets arrived (announce()).
    // Resources.  Operation number i uses flag bank i % kSeqBanks with its own stream and status words.  Small operands
    // are staged in the bank's own area (two halves, alternating between consecutive uses of the bank, which lets the
    // one-shot kernel drop its closing barrier); large ones take a big staging slot in turn; symmetric operands none.
    // Co-residency.  Small operations run with at most small_grid_ blocks, so a full set of banks fits on the GPU
    // together; an operation that wants more waits (on the stream) for everything launched before it, and everything
    // launched after it waits for it.  Those rules depend on the launch sequence only, hence are the same everywhere.
    DeviceReduceResult reduce(const DeviceReduceArgs &a, const std::function<bool()> &abort_requested) override {
        DeviceReduceResult res;
        std::vector<void *> arenas;
        void *mc_view = nullptr;
        bool ring_ok = false;
        {
            std::lock_guard<std::mutex> lock(mu_);
            ring_ok = ring_ready_ && a.world == ring_.size() && (!a.use_nvls || mc_ready_);
            if (ring_ok) {
                if (a.use_nvls) mc_view = reinterpret_cast<void *>(mc_.ptr);
                for (const auto &u : ring_) arenas.push_back(u == self_ ? reinterpret_cast<void *>(arena_ptr_) : reinterpret_cast<void *>(peers_.at(u).ptr));
            }
        }
        if (!ring_ok) {
            if (!a.leased) cancel(a.seq);
            return res;
        }
        cudaSetDevice(device_);

        const size_t esz = wire::dtype_size(a.dtype);
        const int dtype = pccl_dtype(a.dtype), op = pccl_op(a.op), quant = kernel_quant(a.quant);
        const uint64_t bytes = a.count * esz;
        const bool in_place = a.src == a.dst && !a.outer;

        // symmetric memory: the buffer itself is peer-visible, no staging copy for that side
        SymRef ssym, dsym;
        if (!a.outer && quant == 0) {
            std::lock_guard<std::mutex> lock(mu_);
            ssym = symmetric_ref(a.src, bytes);
            dsym = symmetric_ref(a.dst, bytes);
        }
        size_t need = pcclKernelP2PStagingBytes(a.count, dtype, quant);
        if ((ssym.ok != dsym.ok) && need > slot_bytes_) ssym.ok = dsym.ok = false; // mixed and multi-round: stage both sides
        const bool needs_staging = !(ssym.ok && dsym.ok);

        // ---- place in the common order
        std::unique_lock<std::mutex> launch(launch_mu_);
        // (a peer whose sequencer set-up failed must not diverge from the others: whether the log is used is part of the vote)
        const bool sequenced = sequencer_.attached() && (a.leased || a.sequenced);
        if ((a.leased || a.sequenced) && !sequenced) {
            if (!a.leased) cancel(a.seq);
            return res;
        }
        uint64_t index = 0;
        bool queue_turn = false;
        if (sequenced) {
            if (!a.leased) cancel(a.seq); // the Commence queue is not used while the sequencer orders everything
            const bool got = a.leased ? sequencer_.order_of(a.tag, a.lease_k, index) : sequencer_.order_of(~0ull, a.seq, index);
            if (!got) {
                PLOG(ERR) << "same-box sequencer did not answer (a peer died while appending?)";
                return res;
            }
            while (next_launch_ != index) {
                launch_cv_.wait_for(launch, std::chrono::milliseconds(10));
                if (next_launch_ != index && ((abort_requested && abort_requested()) || !sequencer_.attached() || next_launch_ > index)) {
                    res.revoked = a.leased;
                    return res;
                }
            }
        } else {
            launch.unlock();
            std::unique_lock<std::mutex> lock(order_mu_);
            auto &q = order_;
            order_cv_.wait(lock, [&] { return q.empty() || q.front() == a.seq || std::find(q.begin(), q.end(), a.seq) == q.end(); });
            lock.unlock();
            launch.lock();
            queue_turn = true;
        }
        // from here on the turn must be passed on whatever happens
        struct TurnGuard {
            CudaPlane *p;
            bool sequenced, queue_turn, keep_queue;
            uint64_t index, seq;
            std::unique_lock<std::mutex> &launch;
            ~TurnGuard() {
                if (sequenced) {
                    if (!launch.owns_lock()) launch.lock();
                    if (p->next_launch_ == index) p->next_launch_ = index + 1;
                    launch.unlock();
                    p->launch_cv_.notify_all();
                } else if (queue_turn && !keep_queue) {
                    if (launch.owns_lock()) launch.unlock();
                    p->cancel(seq);
                }
            }
        } turn{this, sequenced, queue_turn, false, index, a.seq, launch};

        const int bank_id = sequenced ? static_cast<int>(index % kSeqBanks) : kSeqBanks + static_cast<int>(a.seq % n_slots_);
        Bank &bank = banks_[bank_id];
        while (bank.busy) {
            retire_if_complete(bank_id);
            if (!bank.busy) break;
            launch_cv_.wait_for(launch, std::chrono::milliseconds(1));
            if (abort_requested && abort_requested()) {
                res.revoked = a.leased;
                return res;
            }
        }
        const uint64_t epoch_base = sequenced ? (sequencer_.flag_epoch(index) + 1) * 1024 : (a.seq + 1) * 1024;

        // ---- staging
        const bool small = need <= small_half_ && !a.outer;
        uint64_t stage_off = 0, stage_cap = 0;
        int slot_id = -1;
        bool double_buffered = false;
        if (needs_staging || a.outer) {
            if (sequenced && small) {
                const uint64_t parity = (index / kSeqBanks) & 1;
                stage_off = small_offset(bank_id) + parity * small_half_;
                stage_cap = small_half_;
                double_buffered = true;
            } else {
                slot_id = static_cast<int>(sequenced ? big_count_++ % n_slots_ : a.seq % n_slots_);
                stage_off = slot_offset(slot_id);
                stage_cap = slot_bytes_;
            }
        }
        // rounds: the staging slot bounds how much one launch can move
        uint64_t per_round = a.count;
        if (needs_staging && need > stage_cap) {
            const uint64_t unit = 32 * 4 * a.world * 16; // keeps chunk and quantisation-block alignment
            per_round = (stage_cap - 4096) / (quant ? 2 : esz);
            per_round = std::max<uint64_t>(unit, per_round / unit * unit);
        }
        const uint64_t n_rounds = (a.count + per_round - 1) / per_round;
        // every round gets kRoundFlags barrier values of the operation's 1024 (the pipelined kernel uses one per segment + 2)
        constexpr uint64_t kRoundFlags = 32;
        if (n_rounds > 1024 / kRoundFlags - 1 || (a.outer && n_rounds != 1)) return res;

        // ---- stream dependencies
        bank.busy = true;
        cudaStream_t stream = bank.stream;
        cudaEventRecord(bank.ready, a.user_stream ? static_cast<cudaStream_t>(a.user_stream) : cudaStreamLegacy);
        cudaStreamWaitEvent(stream, bank.ready, 0); // inputs may still be in flight where the caller produced them
        const bool big = !small;
        if (sequenced) {
            if (big) {
                for (int b = 0; b < kSeqBanks; ++b)
                    if (b != bank_id && banks_[b].used) cudaStreamWaitEvent(stream, banks_[b].stop, 0);
                last_big_bank_ = bank_id;
            } else if (last_big_bank_ >= 0 && last_big_bank_ != bank_id) {
                cudaStreamWaitEvent(stream, banks_[last_big_bank_].stop, 0);
            }
        }
        if (slot_id >= 0 && slots_[slot_id].used) cudaStreamWaitEvent(stream, slots_[slot_id].last_use, 0);
        volatile uint32_t *words = reinterpret_cast<volatile uint32_t *>(bank.host_words);
        words[0] = 0; // abort flag
        words[1] = 0; // status
        for (int i = 0; i < 8; ++i) reinterpret_cast<volatile uint64_t *>(bank.host_words + 16)[i] = 0; // phase stamps
        if (a.want_hash) cudaMemsetAsync(bank.hash_dev, 0, 2 * sizeof(uint64_t), stream);
        cudaEventRecord(bank.start, stream);

        int algorithm = 0;
        bool launch_failed = false;
        for (uint64_t r = 0; r < n_rounds && !launch_failed; ++r) {
            const uint64_t off = r * per_round, n = std::min(per_round, a.count - off);
            pcclKernelP2PArgs_t k{};
            for (size_t i = 0; i < arenas.size(); ++i) k.arena[i] = arenas[i];
            k.rank = a.rank;
            k.world = a.world;
            k.flag_slot = static_cast<uint32_t>(bank_id);
            k.flag_base = epoch_base + r * kRoundFlags;
            k.src = static_cast<const uint8_t *>(a.src) + off * esz;
            k.dst = static_cast<uint8_t *>(a.dst) + off * esz;
            k.count = n;
            k.dtype = dtype;
            k.op = op;
            k.quant_algorithm = quant;
            k.src_in_arena = ssym.ok ? 1 : 0;
            k.dst_in_arena = dsym.ok ? 1 : 0;
            k.data_offset = ssym.ok ? ssym.offset : stage_off;
            k.red_offset = dsym.ok ? dsym.offset : stage_off;
            k.data_capacity = needs_staging || a.outer ? stage_cap : n * esz + 512;
            k.max_blocks = small && sequenced ? small_grid_ : max_blocks_;
            k.no_trailing_barrier = double_buffered && !ssym.ok ? 1 : 0;
            k.abort_flag = words;
            k.status = bank.host_words + 1;
            k.stamps = reinterpret_cast<uint64_t *>(bank.host_words + 16);
            k.hash_out = a.want_hash && n_rounds == 1 ? bank.hash_dev : nullptr;
            k.timeout_cycles = timeout_ms_ * static_cast<uint64_t>(clock_khz_);
            if (tune_) { // re-read the tuning knobs on every call (benchmark sweeps)
                if (!(small && sequenced)) k.max_blocks = static_cast<int>(env_u64("PCCL_DEVICE_MAX_BLOCKS", 0));
                k.algorithm = static_cast<int>(env_u64("PCCL_DEVICE_VARIANT", 0));
            }
            int rc;
            // warp-specialised pipeline (copy-in || switch reduction || copy-out) for staged operands of one round; not for
            // in-place operands, whose output may only be written once the whole reduction is known to be complete
            const bool kernel_pipeline = mc_view && !ssym.ok && !dsym.ok && quant == 0 && !in_place &&
                                         n * esz >= pipeline_min_bytes_ && (tune_ ? env_flag("PCCL_KERNEL_PIPELINE", kernel_pipeline_) : kernel_pipeline_);
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
                    rc = kernel_pipeline ? pcclKernelAllReduceNvlsPipelined(&nk, &step, segments, stream) : pcclKernelOuterStepNvls(&nk, &step, stream);
                    algorithm = 3;
                } else {
                    rc = pcclKernelOuterStepP2P(&k, &step, stream);
                    algorithm = 2;
                }
            } else if (mc_view) {
                pcclKernelNvlsArgs_t nk{};
                nk.p2p = k;
                nk.mc_arena = mc_view;
                rc = kernel_pipeline ? pcclKernelAllReduceNvlsPipelined(&nk, nullptr, segments, stream) : pcclKernelAllReduceNvls(&nk, stream);
                algorithm = 3;
            } else {
                rc = pcclKernelAllReduceP2P(&k, &algorithm, stream);
            }
            if (rc != 0) {
                PLOG(ERR) << "all-reduce kernel launch failed: " << cudaGetErrorString(static_cast<cudaError_t>(rc));
                launch_failed = true;
            }
        }
        if (a.want_hash && !launch_failed)
            cudaMemcpyAsync(bank.hash_host, bank.hash_dev, 2 * sizeof(uint64_t), cudaMemcpyDeviceToHost, stream);
        cudaEventRecord(bank.stop, stream);
        bank.used = true;
        if (slot_id >= 0) {
            cudaEventRecord(slots_[slot_id].last_use, stream);
            slots_[slot_id].used = true;
        }
        if (a.user_stream) cudaStreamWaitEvent(static_cast<cudaStream_t>(a.user_stream), bank.stop, 0);

        Flight f;
        f.bank = bank_id;
        f.tag = a.tag;
        f.leased = a.leased;
        f.algorithm = algorithm;
        f.launches = static_cast<uint32_t>(n_rounds);
        f.wire_bytes = quant ? a.count : bytes;
        f.world = a.world;
        f.want_hash = a.want_hash && n_rounds == 1;
        f.destroys_input = (in_place && ssym.ok) || (in_place && n_rounds > 1);
        f.launch_failed = launch_failed;
        f.queue_turn = queue_turn;
        f.seq = a.seq;
        const uint64_t ticket = ++ticket_counter_;
        {
            std::lock_guard<std::mutex> lock(inflight_mu_);
            inflight_.emplace(ticket, f);
        }
        bank.ticket = ticket;
        // pass the turn on: the next operation of the common order may launch now
        if (sequenced) {
            next_launch_ = index + 1;
            launch.unlock();
            launch_cv_.notify_all();
        } else {
            launch.unlock();
            turn.keep_queue = true; // one at a time: the queue entry is removed when this operation has finished
        }
        if (a.enqueue_only && !launch_failed) {
            res.pending = true;
            res.ticket = ticket;
            return res;
        }
        return finish(ticket, abort_requested);
    }

    DeviceReduceResult finish(uint64_t ticket, const std::function<bool()> &abort_requested) override {
        DeviceReduceResult res;
        Flight f;
        {
            std::lock_guard<std::mutex> launch(launch_mu_); // against retire_if_complete()
            std::lock_guard<std::mutex> lock(inflight_mu_);
            const auto it = inflight_.find(ticket);
            if (it == inflight_.end()) return res;
            f = it->second;
            if (f.retired) { // the bank was needed again before the caller came back: the verdict was taken then
                res = f.result;
                inflight_.erase(it);
                return res;
            }
            banks_[f.bank].collecting = true;
        }
        cudaSetDevice(device_);
        Bank &bank = banks_[f.bank];
        volatile uint32_t *words = reinterpret_cast<volatile uint32_t *>(bank.host_words);
        // wait, relaying aborts from the control plane into the kernel
        int polls = 0;
        bool failed = f.launch_failed;
        for (;;) {
            const cudaError_t q = cudaEventQuery(bank.stop);
            if (q == cudaSuccess) break;
            if (q != cudaErrorNotReady) {
                PLOG(ERR) << "device all-reduce failed: " << cudaGetErrorString(q);
                (void) cudaGetLastError();
                failed = true;
                break;
            }
            if ((++polls & 63) == 0 && abort_requested && abort_requested()) {
                words[0] = 1;
                res.aborted = true;
            }
            if (polls > 2000) std::this_thread::sleep_for(std::chrono::microseconds(50));
        }
        res = collect(ticket, f, failed, res.aborted);
        release_bank(f.bank);
        if (f.queue_turn) cancel(f.seq);
        return res;
    }

    bool copy_to_host(void *host_dst, const void *dev_src, size_t bytes) override { return copy(host_dst, dev_src, bytes, cudaMemcpyDeviceToHost); }
    bool copy_to_device(void *dev_dst, const void *host_src, size_t bytes) override { return copy(dev_dst, host_src, bytes, cudaMemcpyHostToDevice); }

    bool hash_device(const void *dev_ptr, size_t bytes, wire::HashType type, uint64_t &out) override {
        if (!cuda_runtime_usable()) return false;
        const PointerInfo info = classify_pointer(dev_ptr);
        int prev = -1;
        cudaGetDevice(&prev);
        if (info.is_device && info.device >= 0) cudaSetDevice(info.device);
        const int rc = pcclKernelHashSync(dev_ptr, bytes, static_cast<int>(type), &out);
        if (prev >= 0) cudaSetDevice(prev);
        if (rc != 0) PLOG(ERR) << "device hash failed: " << cudaGetErrorString(static_cast<cudaError_t>(rc));
        return rc == 0;
    }

    bool hash_device_many(std::vector<HashJob> &jobs, wire::HashType type) override {
        if (!cuda_runtime_usable() || jobs.empty()) return jobs.empty();
        const PointerInfo info = classify_pointer(jobs[0].ptr);
        int prev = -1;
        cudaGetDevice(&prev);
        if (info.is_device && info.device >= 0) cudaSetDevice(info.device);
        std::vector<const void *> ptrs;
        std::vector<size_t> sizes;
        std::vector<uint64_t> out(jobs.size());
        for (const auto &j : jobs) {
            ptrs.push_back(j.ptr);
            sizes.push_back(j.bytes);
        }
        const int rc = pcclKernelHashManySync(ptrs.data(), sizes.data(), jobs.size(), static_cast<int>(type), out.data());
        if (prev >= 0) cudaSetDevice(prev);
        if (rc != 0) {
            PLOG(ERR) << "device hash failed: " << cudaGetErrorString(static_cast<cudaError_t>(rc));
            return false;
        }
        for (size_t i = 0; i < jobs.size(); ++i) jobs[i].result = out[i];
        return true;
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
        // chunks well below the slot size: the two-slot pipeline (stage chunk i+1 while chunk i is read) fills and drains faster
        return arena_ptr_ && !slots_.empty() ? std::min<size_t>(slot_bytes_, size_t(256) << 20) : 0;
    }
    std::unique_lock<std::mutex> exchange_lock() override { return std::unique_lock<std::mutex>(exchange_mu_); }
    size_t exchange_slots() override { return n_slots_; }
    bool export_chunk(size_t slot, const void *src, bool, size_t bytes) override {
        if (bytes > slot_bytes_ || slot >= n_slots_) return false;
        return copy(reinterpret_cast<void *>(arena_ptr_ + slot_offset(slot)), src, bytes, cudaMemcpyDefault);
    }
    bool import_chunk(size_t slot, const Uuid &from, void *dst, bool, size_t bytes) override {
        if (slot >= n_slots_) return false;
        CUdeviceptr base = 0;
        {
            std::lock_guard<std::mutex> lock(mu_);
            const auto it = peers_.find(from);
            if (it == peers_.end() || bytes > slot_bytes_) return false;
            base = it->second.ptr;
        }
        return copy(dst, reinterpret_cast<const void *>(base + slot_offset(slot)), bytes, cudaMemcpyDefault);
    }

    bool import_chunk_begin(size_t slot, const Uuid &from, void *dst, bool, size_t bytes) override {
        if (slot >= n_slots_) return false;
        CUdeviceptr base = 0;
        {
            std::lock_guard<std::mutex> lock(mu_);
            const auto it = peers_.find(from);
            if (it == peers_.end() || bytes > slot_bytes_) return false;
            base = it->second.ptr;
        }
        cudaSetDevice(device_);
        std::lock_guard<std::mutex> lock(copy_mu_);
        if (!util_stream_) RT_TRY(cudaStreamCreateWithFlags(&util_stream_, cudaStreamNonBlocking));
        cudaEvent_t done;
        RT_TRY(cudaEventCreateWithFlags(&done, cudaEventDisableTiming));
        if (cudaMemcpyAsync(dst, reinterpret_cast<const void *>(base + slot_offset(slot)), bytes, cudaMemcpyDefault, util_stream_) != cudaSuccess ||
            cudaEventRecord(done, util_stream_) != cudaSuccess) {
            (void) cudaGetLastError();
            cudaEventDestroy(done);
            return false;
        }
        pending_imports_.push_back(done);
        return true;
    }
    bool import_chunk_finish_oldest() override {
        cudaEvent_t done;
        {
            std::lock_guard<std::mutex> lock(copy_mu_);
            if (pending_imports_.empty()) return false;
            done = pending_imports_.front();
            pending_imports_.pop_front();
        }
        const cudaError_t rc = cudaEventSynchronize(done);
        cudaEventDestroy(done);
        if (rc != cudaSuccess) (void) cudaGetLastError();
        return rc == cudaSuccess;
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
    struct Slot { // a big staging area; consecutive users are ordered through last_use
        cudaEvent_t last_use = nullptr;
        bool used = false;
    };

    size_t slot_offset(uint64_t slot) const { return pcclKernelP2PControlBytes() + small_region_bytes() + slot * slot_bytes_; }
    size_t sym_offset() const { return pcclKernelP2PControlBytes() + small_region_bytes() + size_t(n_slots_) * slot_bytes_; }

    // ---- per-operation bookkeeping -----------------------------------------------------------
    static constexpr int kSeqBanks = 16; // flag banks handed out by position in the common order; the rest serve the Commence-queue order

    struct Bank {
        cudaStream_t stream = nullptr;
        cudaEvent_t start = nullptr, stop = nullptr, ready = nullptr;
        uint32_t *host_words = nullptr; // host-mapped: [0] abort flag, [1] status, [16..32) eight 64-bit phase stamps
        uint64_t *hash_dev = nullptr;   // two raw mset64 sums
        uint64_t *hash_host = nullptr;  // their page-locked landing place
        bool busy = false, used = false, collecting = false; // guarded by launch_mu_
        uint64_t ticket = 0;
    };
    struct Flight {
        int bank = 0;
        uint64_t tag = 0, seq = 0;
        bool leased = false, revoked = false, want_hash = false, destroys_input = false, launch_failed = false, queue_turn = false;
        int algorithm = 0;
        uint32_t launches = 0, world = 1;
        uint64_t wire_bytes = 0;
        bool retired = false;
        DeviceReduceResult result;
    };
    struct SymRef {
        bool ok = false;
        uint64_t offset = 0; // from the start of the arena
    };

    // mu_ held.  An operand counts as symmetric only when the kernels can treat it as such: 16-byte aligned, and its
    // extent rounded up to whole 16-byte packets stays inside ONE allocation of the symmetric heap (the kernels read and
    // write whole packets; a view into the middle of an allocation must not spill into its neighbour).
    SymRef symmetric_ref(const void *ptr, uint64_t bytes) const {
        SymRef r;
        if (!arena_ptr_ || sym_bytes_ == 0 || !ptr) return r;
        const auto p = reinterpret_cast<uintptr_t>(ptr), heap = static_cast<uintptr_t>(arena_ptr_) + sym_offset();
        if (p < heap || p >= heap + sym_bytes_ || (p - static_cast<uintptr_t>(arena_ptr_)) % 16 != 0) return r;
        const uint64_t begin = p - heap, end = begin + round_up(bytes, 16);
        const uint64_t scratch = sym_bytes_ - internal_bytes_;
        if (begin >= scratch) { // the library's own scratch region is one allocation
            r.ok = end <= sym_bytes_;
        } else {
            auto it = sym_allocs_.upper_bound(begin);
            if (it != sym_allocs_.begin()) {
                --it;
                r.ok = begin >= it->first && end <= it->first + it->second;
            }
        }
        r.offset = p - static_cast<uintptr_t>(arena_ptr_);
        return r;
    }

    // mu_ held.  Through the switch every GPU moves about N(1+1/P) bytes per direction (its own stores are echoed back),
    // peer-to-peer two-shot moves 2N(P-1)/P: multicast pays off from nvls_min_world_ peers.
    bool nvls_eligible_locked(const DeviceReduceArgs &a) const {
        return a.quant == wire::QuantAlgo::None && a.world >= nvls_min_world_ && a.count * wire::dtype_size(a.dtype) >= nvls_min_bytes_ &&
               pcclKernelNvlsSupported(pccl_dtype(a.dtype), pccl_op(a.op)) != 0;
    }

    DeviceReduceResult collect(uint64_t ticket, const Flight &f, bool failed, bool aborted) {
        DeviceReduceResult res;
        Bank &bank = banks_[f.bank];
        const uint32_t status = reinterpret_cast<volatile uint32_t *>(bank.host_words)[1];
        if (!failed) cudaEventElapsedTime(&res.device_ms, bank.start, bank.stop);
        res.aborted = aborted;
        res.kernel_launches = f.launches;
        res.plane = f.algorithm == 1 ? pcclExtPlaneP2POneShot : (f.algorithm == 3 ? pcclExtPlaneNvls : pcclExtPlaneP2PTwoShot);
        res.tx_bytes = res.rx_bytes = f.algorithm == 1   ? f.wire_bytes * (f.world - 1)
                                      : f.algorithm == 3 ? f.wire_bytes + f.wire_bytes / f.world
                                                         : 2 * f.wire_bytes * (f.world - 1) / f.world;
        res.ok = !failed && status == 0;
        if (status == 2) PLOG(ERR) << "device all-reduce timed out waiting for a peer";
        if (res.ok && f.want_hash) {
            res.has_hash = true;
            res.hash[0] = bank.hash_host[0];
            res.hash[1] = bank.hash_host[1];
        }
        {
            std::lock_guard<std::mutex> lock(inflight_mu_);
            const auto it = inflight_.find(ticket);
            if (it != inflight_.end()) {
                res.revoked = !res.ok && it->second.revoked;
                inflight_.erase(it);
            }
        }
        res.input_destroyed = !res.ok && f.destroys_input;
        for (int i = 0; i < 8; ++i) res.phase_ns[i] = reinterpret_cast<volatile uint64_t *>(bank.host_words + 16)[i];
        return res;
    }

    void release_bank(int bank_id) {
        {
            std::lock_guard<std::mutex> lock(launch_mu_);
            banks_[bank_id].busy = false;
            banks_[bank_id].collecting = false;
        }
        launch_cv_.notify_all();
    }

    // launch_mu_ held.  A bank whose enqueue-only operation has finished on the GPU but was not collected yet is freed
    // for the next operation; the verdict waits in the flight record.
    void retire_if_complete(int bank_id) {
        Bank &bank = banks_[bank_id];
        if (!bank.busy || bank.collecting || cudaEventQuery(bank.stop) != cudaSuccess) return;
        Flight f;
        {
            std::lock_guard<std::mutex> lock(inflight_mu_);
            const auto it = inflight_.find(bank.ticket);
            if (it == inflight_.end() || it->second.retired) return;
            f = it->second;
        }
        const uint64_t ticket = bank.ticket;
        DeviceReduceResult res = collect(ticket, f, f.launch_failed, false);
        f.retired = true;
        f.result = res;
        {
            std::lock_guard<std::mutex> lock(inflight_mu_);
            inflight_[ticket] = f;
        }
        bank.busy = false;
    }

    size_t small_offset(int bank_id) const { return pcclKernelP2PControlBytes() + size_t(bank_id) * 2 * small_half_; }
    size_t small_region_bytes() const { return round_up(size_t(PCCL_K_FLAG_BANKS) * 2 * small_half_, 2u << 20); }

    // Same-box sequencer of the current ring (see sequencer.hpp): the first member of the ring creates the shared log,
    // the others fetch its descriptor over the rendezvous socket.  Failure just means no leases and Commence-queue order.
    // mu_ held.
    void setup_sequencer(const std::vector<Uuid> &ring) {
        const uint64_t key = ring_key_of(ring);
        void *mem = nullptr;
        int fd = -1;
        if (ring[0] == self_) {
            fd = ::memfd_create("pccl-sequencer", MFD_CLOEXEC);
            if (fd < 0 || ::ftruncate(fd, static_cast<off_t>(Sequencer::bytes())) != 0) {
                if (fd >= 0) ::close(fd);
                return;
            }
            mem = ::mmap(nullptr, Sequencer::bytes(), PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
            if (mem == MAP_FAILED) {
                ::close(fd);
                return;
            }
            Sequencer::format(mem, Sequencer::fresh_base());
            std::lock_guard<std::mutex> lock(seq_srv_mu_);
            seq_fd_ = fd;
            seq_key_ = key;
        } else {
            const auto deadline = clock_type::now() + std::chrono::seconds(10);
            while (fd < 0 && clock_type::now() < deadline) {
                RendezvousReply rep{};
                if (ask(ring[0], kReqSequencer, key, rep, fd) && rep.status == 0 && fd >= 0) break;
                if (fd >= 0) ::close(fd);
                fd = -1;
                if (rep.status == 1) return; // the leader has no sequencer
                std::this_thread::sleep_for(std::chrono::milliseconds(2));
            }
            if (fd < 0) return;
            mem = ::mmap(nullptr, Sequencer::bytes(), PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
            ::close(fd);
            if (mem == MAP_FAILED) return;
        }
        std::lock_guard<std::mutex> lock(launch_mu_);
        seq_mem_ = mem;
        if (!sequencer_.attach(mem)) {
            ::munmap(mem, Sequencer::bytes());
            seq_mem_ = nullptr;
            return;
        }
        next_launch_ = 0;
        PLOG(INFO) << "same-box sequencer attached (flag epochs from " << sequencer_.base() << ")";
    }

    // launch_mu_ held
    void drop_sequencer() {
        sequencer_.detach();
        if (seq_mem_) ::munmap(seq_mem_, Sequencer::bytes());
        seq_mem_ = nullptr;
        std::lock_guard<std::mutex> lock(seq_srv_mu_);
        if (seq_fd_ >= 0) ::close(seq_fd_);
        seq_fd_ = -1;
        seq_key_ = 0;
    }

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
        case wire::QuantAlgo::ZeroPointScale: return pcclQuantZeroPointScale;
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
        arena_bytes_ = round_up(pcclKernelP2PControlBytes() + small_region_bytes() + size_t(n_slots_) * slot_bytes_ + sym_bytes_, gran);
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
        for (auto &sl : slots_) RT_TRY(cudaEventCreateWithFlags(&sl.last_use, cudaEventDisableTiming));
        banks_ = std::vector<Bank>(kSeqBanks + n_slots_);
        for (auto &b : banks_) {
            RT_TRY(cudaStreamCreateWithFlags(&b.stream, cudaStreamNonBlocking));
            RT_TRY(cudaEventCreate(&b.start));
            RT_TRY(cudaEventCreate(&b.stop));
            RT_TRY(cudaEventCreateWithFlags(&b.ready, cudaEventDisableTiming));
            RT_TRY(cudaHostAlloc(reinterpret_cast<void **>(&b.host_words), 256, cudaHostAllocMapped));
            std::memset(b.host_words, 0, 256);
            RT_TRY(cudaMalloc(reinterpret_cast<void **>(&b.hash_dev), 2 * sizeof(uint64_t)));
            RT_TRY(cudaHostAlloc(reinterpret_cast<void **>(&b.hash_host), 2 * sizeof(uint64_t), cudaHostAllocDefault));
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
            // The abstract socket namespace is open to every local user: only processes of this user get descriptors of the
            // arena (whoever holds one can read and write this peer's tensors).
            ucred cred{};
            socklen_t cred_len = sizeof cred;
            if (::getsockopt(c, SOL_SOCKET, SO_PEERCRED, &cred, &cred_len) != 0 || cred.uid != ::geteuid()) {
                PLOG(WARN) << "rendezvous request from another user refused";
                ::close(c);
                continue;
            }
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
            } else if (got && req.kind == kReqSequencer) {
                std::lock_guard<std::mutex> lock(seq_srv_mu_);
                if (!use_sequencer_) {
                    rep.status = 1;
                } else if (seq_fd_ >= 0 && seq_key_ == req.ring_key) {
                    fd = ::dup(seq_fd_);
                    rep.status = fd >= 0 ? 0 : 1;
                } else {
                    rep.status = 2; // not (yet) created for that ring: ask again
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

    // launch order, flag banks, operations in flight
    std::mutex launch_mu_;
    std::condition_variable launch_cv_;
    Sequencer sequencer_;
    void *seq_mem_ = nullptr;
    std::mutex seq_srv_mu_; // descriptor of the log this peer created, shared with the server thread
    int seq_fd_ = -1;
    uint64_t seq_key_ = 0;
    bool use_sequencer_ = true;
    uint64_t next_launch_ = 0, big_count_ = 0;
    int last_big_bank_ = -1;
    int small_grid_ = 8;
    size_t small_half_ = 1u << 20;
    std::vector<Bank> banks_;
    std::mutex inflight_mu_;
    std::map<uint64_t, Flight> inflight_;
    std::atomic<uint64_t> ticket_counter_{0};
    std::mutex exchange_mu_;

    std::mutex order_mu_;
    std::condition_variable order_cv_;
    std::deque<uint64_t> order_; // announced device operations in Commence arrival order

    std::mutex copy_mu_;
    cudaStream_t util_stream_ = nullptr;
    std::deque<cudaEvent_t> pending_imports_; // guarded by copy_mu_

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
    // Hot path of small reductions: the same few pointers are asked about again and again.  Only positive answers are
    // remembered (unified addressing keeps device and host ranges apart, so a device address stays a device address of
    // that GPU for as long as it is valid at all).
    struct Known {
        const void *ptr;
        int device;
    };
    thread_local Known known[8] = {};
    thread_local unsigned next_slot = 0;
    for (const Known &k : known)
        if (k.ptr == p) {
            info.is_device = true;
            info.device = k.device;
            return info;
        }
    cudaPointerAttributes attr{};
    if (cudaPointerGetAttributes(&attr, p) != cudaSuccess) {
        (void) cudaGetLastError();
        return info;
    }
    if (attr.type == cudaMemoryTypeDevice || attr.type == cudaMemoryTypeManaged) {
        info.is_device = true;
        info.device = attr.device;
        known[next_slot++ % 8] = Known{p, attr.device};
    }
    return info;
}

std::unique_ptr<Plane> Plane::create() { return std::make_unique<CudaPlane>(); }

} // namespace pccl::device
#endif // PCCL_NO_CUDA
