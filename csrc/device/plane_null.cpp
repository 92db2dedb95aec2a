// Inert device plane used when the library is built without CUDA (PCCL_NO_CUDA).
#include <cstdlib>

#include "plane.hpp"

namespace pccl::device {

#ifdef PCCL_NO_CUDA
namespace {
class NullPlane final : public Plane {
    std::mutex mu_, fuse_use_mu_;

public:
    void start(const Uuid &) override {}
    void stop() override {}
    void on_ring_changed(const std::vector<Uuid> &) override {}
    uint32_t capabilities(const DeviceReduceArgs &) override { return 0; }
    void announce(uint64_t) override {}
    void cancel(uint64_t) override {}
    DeviceReduceResult reduce(const DeviceReduceArgs &, const std::function<bool()> &) override { return {}; }
    bool lease_ready() override { return false; }
    bool multicast_ready() override { return false; }
    bool nvls_eligible(const DeviceReduceArgs &) override { return false; }
    void revoke_leased(uint64_t, bool) override {}
    DeviceReduceResult finish(uint64_t, const std::function<bool()> &) override { return {}; }
    bool copy_to_host(void *, const void *, size_t) override { return false; }
    bool copy_to_device(void *, const void *, size_t) override { return false; }
    bool hash_device(const void *, size_t, wire::HashType, uint64_t &) override { return false; }
    bool hash_device_many(std::vector<HashJob> &, wire::HashType) override { return false; }
    size_t exchange_slots() override { return 0; }
    bool import_chunk_begin(size_t, const Uuid &, void *, bool, size_t) override { return false; }
    bool import_chunk_finish_oldest() override { return false; }
    void *fuse_buffer(size_t) override { return nullptr; }
    std::unique_lock<std::mutex> fuse_lock() override { return std::unique_lock<std::mutex>(fuse_use_mu_); }
    bool gather(const std::vector<Part> &, void *, bool) override { return false; }
    bool peer_mapped(const Uuid &) override { return false; }
    size_t exchange_bytes() override { return 0; }
    std::unique_lock<std::mutex> exchange_lock() override { return std::unique_lock<std::mutex>(mu_); }
    bool export_chunk(size_t, const void *, bool, size_t) override { return false; }
    bool import_chunk(size_t, const Uuid &, void *, bool, size_t) override { return false; }
    bool pseudo_grad(float *, const float *, const float *, size_t) override { return false; }
    bool outer_nesterov(const OuterStepArgs &, const float *, size_t) override { return false; }
    void *host_alloc(size_t bytes) override { return std::malloc(bytes); }
    void host_free(void *ptr) override { std::free(ptr); }
    void *symmetric_alloc(size_t) override { return nullptr; }
    bool symmetric_free(void *) override { return false; }
};
} // namespace

PointerInfo classify_pointer(const void *) { return {}; }
bool cuda_runtime_usable() { return false; }
std::unique_ptr<Plane> Plane::create() { return std::make_unique<NullPlane>(); }
#endif

} // namespace pccl::device
