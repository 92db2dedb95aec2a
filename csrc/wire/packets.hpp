// All CCoIP packets.  Field order and widths follow the reference byte for byte so that a peer
// or master built from the reference could interoperate (reference: ccoip/internal/ccoip_packets.hpp:31-417,
// ccoip/src/cpp/ccoip_packets.cpp:21-625).  Packet ids are per direction.
//
// Extension packets of this build use ids >= 64 and are only ever sent to endpoints that
// announced support, so the base protocol stays untouched.
#pragma once
#include <optional>
#include <string>
#include <vector>

#include "../util/uuid.hpp"
#include "buffer.hpp"
#include "pccl.h"

namespace pccl::wire {

using PacketId = uint16_t;

// Data types as numbered on the wire (reference: ccoip/internal/ccoip_types.hpp:8-21).  NOTE: differs
// from pcclDataType_t numbering.
enum class DType : uint8_t { U8 = 0, I8 = 1, U16 = 2, U32 = 3, I16 = 4, I32 = 5, U64 = 6, I64 = 7, F16 = 8, BF16 = 9, F32 = 10, F64 = 11, F8E4M3 = 12 };
enum class RedOp : uint8_t { Set = 0, Sum = 1, Avg = 2, Prod = 3, Max = 4, Min = 5 };
enum class HashType : uint8_t { Simple = 0, Crc32 = 1, Mset64 = 2 }; // 2: extension, see kernels/mset_formula.hpp
enum class SyncStrategy : uint8_t { EnforcePopular = 0, RxOnly = 1, TxOnly = 2 };
enum class QuantAlgo : uint8_t { None = 0, MinMax = 1, ZeroPointScale = 2, BlockMinMax = 3, BlockFp8 = 4 };

size_t dtype_size(DType t);
const char *dtype_name(DType t);

void put_sockaddr(Writer &w, const ccoip_socket_address_t &a);
ccoip_socket_address_t get_sockaddr(Reader &r);
void put_uuid(Writer &w, const Uuid &u);
Uuid get_uuid(Reader &r);

#define PCCL_PACKET(ID)                                                                                                \
    static constexpr PacketId id = ID;                                                                                 \
    void write(Writer &w) const;                                                                                       \
    bool read(Reader &r);

#define PCCL_EMPTY_PACKET(ID)                                                                                          \
    static constexpr PacketId id = ID;                                                                                 \
    void write(Writer &) const {}                                                                                      \
    bool read(Reader &) { return true; }

// ---------------------------------------------------------------- client -> master
struct C2MSessionRegistration {
    PCCL_PACKET(1)
    uint32_t peer_group = 0;
    bool explicit_addresses = false;
    ccoip_socket_address_t adv_p2p{}, adv_ss{}, adv_bm{};
    uint16_t p2p_port = 0, ss_port = 0, bm_port = 0;
};
struct C2MEstablishP2P {
    PCCL_PACKET(2)
    bool accept_new_peers = false;
};
struct C2MP2PEstablished {
    PCCL_PACKET(3)
    bool success = false;
    std::vector<Uuid> failed_peers;
};
struct C2MCheckPeersPending {
    PCCL_EMPTY_PACKET(4)
};
struct C2MOptimizeTopology {
    PCCL_EMPTY_PACKET(5)
};
struct C2MReportBandwidth {
    PCCL_PACKET(6)
    Uuid to_peer;
    double mbits_per_second = 0;
};
struct C2MOptimizeWorkComplete {
    PCCL_EMPTY_PACKET(7)
};
struct SharedStateEntry {
    std::string key;
    uint64_t hash = 0;
    HashType hash_type = HashType::Simple;
    uint64_t num_elements = 0;
    DType dtype = DType::U8;
    bool allow_content_inequality = false;
    bool operator==(const SharedStateEntry &o) const {
        return key == o.key && hash == o.hash && hash_type == o.hash_type && num_elements == o.num_elements &&
               dtype == o.dtype && allow_content_inequality == o.allow_content_inequality;
    }
};
struct C2MSyncSharedState {
    PCCL_PACKET(8)
    uint64_t revision = 0;
    SyncStrategy strategy = SyncStrategy::EnforcePopular;
    std::vector<SharedStateEntry> entries;
};
struct C2MDistSharedStateComplete {
    PCCL_EMPTY_PACKET(9)
};
struct C2MCollectiveInitiate {
    PCCL_PACKET(10)
    uint64_t tag = 0;
    uint64_t count = 0;
    DType dtype = DType::F32;
    RedOp op = RedOp::Sum;
};
struct C2MCollectiveComplete {
    PCCL_PACKET(11)
    uint64_t tag = 0;
    bool was_aborted = false;
};

// ---------------------------------------------------------------- master -> client
struct M2CSessionRegistrationResponse {
    PCCL_PACKET(1)
    bool accepted = false;
    Uuid assigned_uuid;
};
struct PeerEndpoint {
    ccoip_socket_address_t p2p_addr{};
    Uuid uuid;
};
struct M2CP2PConnectionInfo {
    PCCL_PACKET(2)
    bool unchanged = false;
    uint64_t global_world_size = 0, local_world_size = 0, num_peer_groups = 0, largest_group_size = 0;
    std::vector<PeerEndpoint> peers; // only on the wire when !unchanged
};
struct M2CP2PEstablished {
    PCCL_PACKET(3)
    bool success = false;
    std::vector<Uuid> ring_order;
};
struct M2CPeersPendingResponse {
    PCCL_PACKET(4)
    bool pending = false;
};
struct BenchmarkRequest {
    Uuid from, to;
    ccoip_socket_address_t endpoint{};
};
struct M2COptimizeTopologyResponse {
    PCCL_PACKET(5)
    std::vector<BenchmarkRequest> requests;
};
struct M2COptimizeTopologyComplete {
    PCCL_PACKET(6)
    bool success = false;
    std::vector<Uuid> ring_order;
};
struct M2CSyncSharedState {
    PCCL_PACKET(7)
    bool is_outdated = false;
    ccoip_socket_address_t distributor{};
    std::vector<std::string> outdated_keys;
    std::vector<uint64_t> expected_hashes;
    std::vector<HashType> expected_hash_types;
};
struct M2CSyncSharedStateComplete {
    PCCL_EMPTY_PACKET(8)
};
struct M2CCollectiveCommence {
    PCCL_PACKET(9)
    uint64_t tag = 0;
    uint64_t seq_nr = 0;
};
struct M2CCollectiveComplete {
    PCCL_PACKET(10)
    uint64_t tag = 0;
};
struct M2CCollectiveAbort {
    PCCL_PACKET(11)
    uint64_t tag = 0;
    bool aborted = false;
};
// Extension (never sent to a group that contains a peer without lease support, so reference peers do not see it): the
// lease under which a group ran the collectives of `tag` (or of every tag, `all`) without voting has ended.  `peer_lost`:
// because a member of the group is gone -- the ring must be re-established before the next operation.
struct M2CLeaseRevoked {
    PCCL_PACKET(12)
    uint64_t tag = 0;
    bool all = false;
    bool peer_lost = false;
};

// ---------------------------------------------------------------- peer <-> peer
struct P2PHello {
    PCCL_PACKET(1)
    Uuid uuid;
    uint32_t connection_nr = 0;
};
struct P2PHelloAck {
    PCCL_EMPTY_PACKET(2)
};
enum class DequantMetaType : uint8_t { MinMax = 0, ZeroPointScale = 1 };
struct P2PDequantMeta {
    PCCL_PACKET(3)
    uint64_t tag = 0;
    DequantMetaType meta_type = DequantMetaType::MinMax;
    // MinMax: dtype a = type of min/max, blob a = min, blob b = max.
    // ZeroPointScale: dtype a = zero-point type, blob a = zero point, dtype b = scale type, blob b = scale.
    DType dtype_a = DType::F32, dtype_b = DType::F32;
    std::vector<uint8_t> blob_a, blob_b;
    size_t wire_size() const;
};

// Extension (never sent unless the hierarchy is switched on, see Client::build_hierarchy): peers of one ring tell each other
// which host they run on and where they can be dialled, travelling around the ring like an all-gather.  A peer contributes its
// own host identity and the endpoint of its ring SUCCESSOR (the only endpoints a peer learns from the master are its
// neighbours'), so after world-1 hops everybody knows everybody's host and address.
struct DirectoryEntry {
    Uuid uuid;
    uint64_t host_id = 0;      // valid when has_host
    bool has_host = false, has_endpoint = false;
    ccoip_socket_address_t endpoint{}; // valid when has_endpoint
};
struct P2PDirectory {
    PCCL_PACKET(5)
    std::vector<DirectoryEntry> entries;
};

// ---------------------------------------------------------------- shared-state server
struct C2SRequestSharedState {
    PCCL_PACKET(1)
    std::vector<std::string> keys;
};
enum class SharedStateStatus : uint8_t { Success = 1, NotDistributed = 2, NotInDistributionMode = 3 };
struct SharedStateSlice {
    std::string key;
    uint64_t size_bytes = 0;
};
struct S2CSharedStateResponse {
    PCCL_PACKET(1)
    SharedStateStatus status = SharedStateStatus::Success;
    uint64_t revision = 0;
    std::vector<SharedStateSlice> entries; // raw tensor bytes follow the frame, in this order
};

// Extension (ids >= 2 do not exist in the reference): shared state over NVLink.  The requester
// names itself; a distributor that has the requester's arena mapped answers with p2p = true and
// then announces every staged chunk with S2CChunkReady, the requester copies it out of the
// distributor's arena and answers C2SChunkAck.  Otherwise raw bytes follow as in the base protocol.
struct C2SRequestSharedStateP2P {
    PCCL_PACKET(2)
    Uuid requester;
    std::vector<std::string> keys;
};
struct S2CSharedStateResponseP2P {
    PCCL_PACKET(2)
    SharedStateStatus status = SharedStateStatus::Success;
    uint64_t revision = 0;
    bool p2p = false;
    Uuid distributor;
    uint64_t chunk_bytes = 0;
    std::vector<SharedStateSlice> entries;
};
struct S2CChunkReady {
    PCCL_EMPTY_PACKET(3)
};
struct C2SChunkAck {
    PCCL_EMPTY_PACKET(3)
};

// ---------------------------------------------------------------- bandwidth benchmark
struct C2BHello {
    PCCL_PACKET(1)
    Uuid uuid;
};
struct B2CBusy {
    PCCL_PACKET(1)
    bool busy = false;
};

// Control frame = [u64 BE length = 2 + payload][u16 BE id][payload]
template <typename P> std::vector<uint8_t> encode_frame(const P &pkt) {
    Writer body;
    pkt.write(body);
    Writer w(10 + body.size());
    w.put<uint64_t>(2 + body.size());
    w.put<uint16_t>(P::id);
    w.put_bytes(body.bytes().data(), body.size());
    return std::move(w.bytes());
}

constexpr uint64_t kMaxControlFrame = 64ull << 20; // rx cap for control frames
constexpr uint64_t kMaxP2PFrame = 1ull << 30;      // rx cap for multiplexed data frames

} // namespace pccl::wire
