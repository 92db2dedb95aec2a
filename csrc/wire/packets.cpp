#include "packets.hpp"

namespace pccl::wire {

size_t dtype_size(DType t) {
    switch (t) {
    case DType::U8:
    case DType::I8:
    case DType::F8E4M3: return 1;
    case DType::U16:
    case DType::I16:
    case DType::F16:
    case DType::BF16: return 2;
    case DType::U32:
    case DType::I32:
    case DType::F32: return 4;
    case DType::U64:
    case DType::I64:
    case DType::F64: return 8;
    }
    return 0;
}

const char *dtype_name(DType t) {
    static const char *names[] = {"u8", "i8", "u16", "u32", "i16", "i32", "u64", "i64", "f16", "bf16", "f32", "f64", "f8e4m3"};
    const auto i = static_cast<size_t>(t);
    return i < sizeof(names) / sizeof(names[0]) ? names[i] : "?";
}

void put_sockaddr(Writer &w, const ccoip_socket_address_t &a) {
    const bool v4 = a.inet.protocol == inetIPv4;
    w.put<bool>(v4);
    if (v4)
        w.put_bytes(a.inet.ipv4.data, 4);
    else
        w.put_bytes(a.inet.ipv6.data, 16);
    w.put<uint16_t>(a.port);
}

ccoip_socket_address_t get_sockaddr(Reader &r) {
    ccoip_socket_address_t a{};
    if (r.get<bool>()) {
        a.inet.protocol = inetIPv4;
        r.get_bytes(a.inet.ipv4.data, 4);
    } else {
        a.inet.protocol = inetIPv6;
        r.get_bytes(a.inet.ipv6.data, 16);
    }
    a.port = r.get<uint16_t>();
    return a;
}

void put_uuid(Writer &w, const Uuid &u) { w.put_bytes(u.data.data(), 16); }
Uuid get_uuid(Reader &r) {
    Uuid u;
    r.get_bytes(u.data.data(), 16);
    return u;
}

static void put_uuid_list(Writer &w, const std::vector<Uuid> &v) {
    w.put<uint64_t>(v.size());
    for (const auto &u : v) put_uuid(w, u);
}
static bool get_uuid_list(Reader &r, std::vector<Uuid> &v) {
    const uint64_t n = r.get<uint64_t>();
    if (!r.plausible_count(n, 16)) return false;
    v.clear();
    v.reserve(n);
    for (uint64_t i = 0; i < n; ++i) v.push_back(get_uuid(r));
    return r.ok();
}

// ---- C2M
void C2MSessionRegistration::write(Writer &w) const {
    w.put<uint32_t>(peer_group);
    w.put<bool>(explicit_addresses);
    if (explicit_addresses) {
        put_sockaddr(w, adv_p2p);
        put_sockaddr(w, adv_ss);
        put_sockaddr(w, adv_bm);
    } else {
        w.put<uint16_t>(p2p_port);
        w.put<uint16_t>(ss_port);
        w.put<uint16_t>(bm_port);
    }
}
bool C2MSessionRegistration::read(Reader &r) {
    peer_group = r.get<uint32_t>();
    explicit_addresses = r.get<bool>();
    if (explicit_addresses) {
        adv_p2p = get_sockaddr(r);
        adv_ss = get_sockaddr(r);
        adv_bm = get_sockaddr(r);
        p2p_port = adv_p2p.port;
        ss_port = adv_ss.port;
        bm_port = adv_bm.port;
    } else {
        p2p_port = r.get<uint16_t>();
        ss_port = r.get<uint16_t>();
        bm_port = r.get<uint16_t>();
    }
    return r.ok();
}

void C2MEstablishP2P::write(Writer &w) const { w.put<bool>(accept_new_peers); }
bool C2MEstablishP2P::read(Reader &r) {
    accept_new_peers = r.get<bool>();
    return r.ok();
}

void C2MP2PEstablished::write(Writer &w) const {
    w.put<bool>(success);
    put_uuid_list(w, failed_peers);
}
bool C2MP2PEstablished::read(Reader &r) {
    success = r.get<bool>();
    return get_uuid_list(r, failed_peers);
}

void C2MReportBandwidth::write(Writer &w) const {
    put_uuid(w, to_peer);
    w.put<double>(mbits_per_second);
}
bool C2MReportBandwidth::read(Reader &r) {
    to_peer = get_uuid(r);
    mbits_per_second = r.get<double>();
    return r.ok();
}

void C2MSyncSharedState::write(Writer &w) const {
    w.put<uint64_t>(revision);
    w.put<uint8_t>(static_cast<uint8_t>(strategy));
    w.put<uint64_t>(entries.size());
    for (const auto &e : entries) {
        w.put_string(e.key);
        w.put<uint64_t>(e.hash);
        w.put<uint8_t>(static_cast<uint8_t>(e.hash_type));
        w.put<uint64_t>(e.num_elements);
        w.put<uint8_t>(static_cast<uint8_t>(e.dtype));
        w.put<bool>(e.allow_content_inequality);
    }
}
bool C2MSyncSharedState::read(Reader &r) {
    revision = r.get<uint64_t>();
    strategy = static_cast<SyncStrategy>(r.get<uint8_t>());
    const uint64_t n = r.get<uint64_t>();
    if (!r.plausible_count(n, 8 + 8 + 1 + 8 + 1 + 1)) return false;
    entries.clear();
    for (uint64_t i = 0; i < n && r.ok(); ++i) {
        SharedStateEntry e;
        e.key = r.get_string();
        e.hash = r.get<uint64_t>();
        e.hash_type = static_cast<HashType>(r.get<uint8_t>());
        e.num_elements = r.get<uint64_t>();
        e.dtype = static_cast<DType>(r.get<uint8_t>());
        e.allow_content_inequality = r.get<bool>();
        entries.push_back(std::move(e));
    }
    return r.ok();
}

void C2MCollectiveInitiate::write(Writer &w) const {
    w.put<uint64_t>(tag);
    w.put<uint64_t>(count);
    w.put<uint8_t>(static_cast<uint8_t>(dtype));
    w.put<uint8_t>(static_cast<uint8_t>(op));
}
bool C2MCollectiveInitiate::read(Reader &r) {
    tag = r.get<uint64_t>();
    count = r.get<uint64_t>();
    dtype = static_cast<DType>(r.get<uint8_t>());
    op = static_cast<RedOp>(r.get<uint8_t>());
    return r.ok();
}

void C2MCollectiveComplete::write(Writer &w) const {
    w.put<uint64_t>(tag);
    w.put<bool>(was_aborted);
}
bool C2MCollectiveComplete::read(Reader &r) {
    tag = r.get<uint64_t>();
    was_aborted = r.get<bool>();
    return r.ok();
}

// ---- M2C
void M2CSessionRegistrationResponse::write(Writer &w) const {
    w.put<bool>(accepted);
    put_uuid(w, assigned_uuid);
}
bool M2CSessionRegistrationResponse::read(Reader &r) {
    accepted = r.get<bool>();
    assigned_uuid = get_uuid(r);
    return r.ok();
}

void M2CP2PConnectionInfo::write(Writer &w) const {
    w.put<bool>(unchanged);
    w.put<uint64_t>(global_world_size);
    w.put<uint64_t>(local_world_size);
    w.put<uint64_t>(num_peer_groups);
    w.put<uint64_t>(largest_group_size);
    if (unchanged) return;
    w.put<uint64_t>(peers.size());
    for (const auto &p : peers) {
        put_sockaddr(w, p.p2p_addr);
        put_uuid(w, p.uuid);
    }
}
bool M2CP2PConnectionInfo::read(Reader &r) {
    unchanged = r.get<bool>();
    global_world_size = r.get<uint64_t>();
    local_world_size = r.get<uint64_t>();
    num_peer_groups = r.get<uint64_t>();
    largest_group_size = r.get<uint64_t>();
    peers.clear();
    if (unchanged) return r.ok();
    const uint64_t n = r.get<uint64_t>();
    if (!r.plausible_count(n, 1 + 4 + 2 + 16)) return false;
    for (uint64_t i = 0; i < n && r.ok(); ++i) {
        PeerEndpoint p;
        p.p2p_addr = get_sockaddr(r);
        p.uuid = get_uuid(r);
        peers.push_back(p);
    }
    return r.ok();
}

void M2CP2PEstablished::write(Writer &w) const {
    w.put<bool>(success);
    put_uuid_list(w, ring_order);
}
bool M2CP2PEstablished::read(Reader &r) {
    success = r.get<bool>();
    return get_uuid_list(r, ring_order);
}

void M2CPeersPendingResponse::write(Writer &w) const { w.put<bool>(pending); }
bool M2CPeersPendingResponse::read(Reader &r) {
    pending = r.get<bool>();
    return r.ok();
}

void M2COptimizeTopologyResponse::write(Writer &w) const {
    w.put<uint64_t>(requests.size());
    for (const auto &q : requests) {
        put_uuid(w, q.from);
        put_uuid(w, q.to);
        put_sockaddr(w, q.endpoint);
    }
}
bool M2COptimizeTopologyResponse::read(Reader &r) {
    const uint64_t n = r.get<uint64_t>();
    if (!r.plausible_count(n, 16 + 16 + 1 + 4 + 2)) return false;
    requests.clear();
    for (uint64_t i = 0; i < n && r.ok(); ++i) {
        BenchmarkRequest q;
        q.from = get_uuid(r);
        q.to = get_uuid(r);
        q.endpoint = get_sockaddr(r);
        requests.push_back(q);
    }
    return r.ok();
}

void M2COptimizeTopologyComplete::write(Writer &w) const {
    w.put<bool>(success);
    put_uuid_list(w, ring_order);
}
bool M2COptimizeTopologyComplete::read(Reader &r) {
    success = r.get<bool>();
    return get_uuid_list(r, ring_order);
}

void M2CSyncSharedState::write(Writer &w) const {
    w.put<bool>(is_outdated);
    put_sockaddr(w, distributor);
    w.put<uint64_t>(outdated_keys.size());
    for (const auto &k : outdated_keys) w.put_string(k);
    for (const auto h : expected_hashes) w.put<uint64_t>(h);
    for (const auto t : expected_hash_types) w.put<uint8_t>(static_cast<uint8_t>(t));
}
bool M2CSyncSharedState::read(Reader &r) {
    is_outdated = r.get<bool>();
    distributor = get_sockaddr(r);
    const uint64_t n = r.get<uint64_t>();
    if (!r.plausible_count(n, 8 + 8 + 1)) return false;
    outdated_keys.clear();
    expected_hashes.clear();
    expected_hash_types.clear();
    for (uint64_t i = 0; i < n && r.ok(); ++i) outdated_keys.push_back(r.get_string());
    for (uint64_t i = 0; i < n && r.ok(); ++i) expected_hashes.push_back(r.get<uint64_t>());
    for (uint64_t i = 0; i < n && r.ok(); ++i) expected_hash_types.push_back(static_cast<HashType>(r.get<uint8_t>()));
    return r.ok();
}

void M2CCollectiveCommence::write(Writer &w) const {
    w.put<uint64_t>(tag);
    w.put<uint64_t>(seq_nr);
}
bool M2CCollectiveCommence::read(Reader &r) {
    tag = r.get<uint64_t>();
    seq_nr = r.get<uint64_t>();
    return r.ok();
}
void M2CCollectiveComplete::write(Writer &w) const { w.put<uint64_t>(tag); }
bool M2CCollectiveComplete::read(Reader &r) {
    tag = r.get<uint64_t>();
    return r.ok();
}
void M2CCollectiveAbort::write(Writer &w) const {
    w.put<uint64_t>(tag);
    w.put<bool>(aborted);
}
bool M2CCollectiveAbort::read(Reader &r) {
    tag = r.get<uint64_t>();
    aborted = r.get<bool>();
    return r.ok();
}

void M2CLeaseRevoked::write(Writer &w) const {
    w.put<uint64_t>(tag);
    w.put<bool>(all);
    w.put<bool>(peer_lost);
}
bool M2CLeaseRevoked::read(Reader &r) {
    tag = r.get<uint64_t>();
    all = r.get<bool>();
    peer_lost = r.get<bool>();
    return r.ok();
}

void P2PDirectory::write(Writer &w) const {
    w.put<uint64_t>(entries.size());
    for (const auto &e : entries) {
        put_uuid(w, e.uuid);
        w.put<uint64_t>(e.host_id);
        w.put<bool>(e.has_host);
        w.put<bool>(e.has_endpoint);
        put_sockaddr(w, e.endpoint);
    }
}
bool P2PDirectory::read(Reader &r) {
    const uint64_t n = r.get<uint64_t>();
    if (n > 4096) return false;
    entries.clear();
    for (uint64_t i = 0; i < n && r.ok(); ++i) {
        DirectoryEntry e;
        e.uuid = get_uuid(r);
        e.host_id = r.get<uint64_t>();
        e.has_host = r.get<bool>();
        e.has_endpoint = r.get<bool>();
        e.endpoint = get_sockaddr(r);
        entries.push_back(e);
    }
    return r.ok();
}

// ---- P2P
void P2PHello::write(Writer &w) const {
    put_uuid(w, uuid);
    w.put<uint32_t>(connection_nr);
}
bool P2PHello::read(Reader &r) {
    uuid = get_uuid(r);
    connection_nr = r.get<uint32_t>();
    return r.ok();
}

void P2PDequantMeta::write(Writer &w) const {
    w.put<uint64_t>(tag);
    w.put<uint8_t>(static_cast<uint8_t>(meta_type));
    w.put<uint8_t>(static_cast<uint8_t>(dtype_a));
    if (meta_type == DequantMetaType::MinMax) {
        w.put_bytes(blob_a.data(), blob_a.size());
        w.put_bytes(blob_b.data(), blob_b.size());
    } else {
        w.put_bytes(blob_a.data(), blob_a.size());
        w.put<uint8_t>(static_cast<uint8_t>(dtype_b));
        w.put_bytes(blob_b.data(), blob_b.size());
    }
}
bool P2PDequantMeta::read(Reader &r) {
    tag = r.get<uint64_t>();
    meta_type = static_cast<DequantMetaType>(r.get<uint8_t>());
    dtype_a = static_cast<DType>(r.get<uint8_t>());
    if (meta_type == DequantMetaType::MinMax) {
        const size_t n = dtype_size(dtype_a);
        if (n == 0) return false;
        blob_a.resize(n);
        blob_b.resize(n);
        r.get_bytes(blob_a.data(), n);
        r.get_bytes(blob_b.data(), n);
        dtype_b = dtype_a;
    } else if (meta_type == DequantMetaType::ZeroPointScale) {
        const size_t na = dtype_size(dtype_a);
        if (na == 0) return false;
        blob_a.resize(na);
        r.get_bytes(blob_a.data(), na);
        dtype_b = static_cast<DType>(r.get<uint8_t>());
        const size_t nb = dtype_size(dtype_b);
        if (nb == 0) return false;
        blob_b.resize(nb);
        r.get_bytes(blob_b.data(), nb);
    } else {
        return false;
    }
    return r.ok();
}
size_t P2PDequantMeta::wire_size() const {
    Writer w;
    write(w);
    return w.size();
}

// ---- shared-state server
void C2SRequestSharedState::write(Writer &w) const {
    w.put<uint64_t>(keys.size());
    for (const auto &k : keys) w.put_string(k);
}
bool C2SRequestSharedState::read(Reader &r) {
    const uint64_t n = r.get<uint64_t>();
    if (!r.plausible_count(n, 8)) return false;
    keys.clear();
    for (uint64_t i = 0; i < n && r.ok(); ++i) keys.push_back(r.get_string());
    return r.ok();
}
void S2CSharedStateResponse::write(Writer &w) const {
    w.put<uint8_t>(static_cast<uint8_t>(status));
    w.put<uint64_t>(revision);
    w.put<uint64_t>(entries.size());
    for (const auto &e : entries) {
        w.put_string(e.key);
        w.put<uint64_t>(e.size_bytes);
    }
}
bool S2CSharedStateResponse::read(Reader &r) {
    status = static_cast<SharedStateStatus>(r.get<uint8_t>());
    revision = r.get<uint64_t>();
    const uint64_t n = r.get<uint64_t>();
    if (!r.plausible_count(n, 16)) return false;
    entries.clear();
    for (uint64_t i = 0; i < n && r.ok(); ++i) {
        SharedStateSlice s;
        s.key = r.get_string();
        s.size_bytes = r.get<uint64_t>();
        entries.push_back(std::move(s));
    }
    return r.ok();
}

void C2SRequestSharedStateP2P::write(Writer &w) const {
    put_uuid(w, requester);
    w.put<uint64_t>(keys.size());
    for (const auto &k : keys) w.put_string(k);
}
bool C2SRequestSharedStateP2P::read(Reader &r) {
    requester = get_uuid(r);
    const uint64_t n = r.get<uint64_t>();
    if (!r.plausible_count(n, 8)) return false;
    keys.clear();
    for (uint64_t i = 0; i < n && r.ok(); ++i) keys.push_back(r.get_string());
    return r.ok();
}
void S2CSharedStateResponseP2P::write(Writer &w) const {
    w.put<uint8_t>(static_cast<uint8_t>(status));
    w.put<uint64_t>(revision);
    w.put<bool>(p2p);
    put_uuid(w, distributor);
    w.put<uint64_t>(chunk_bytes);
    w.put<uint64_t>(entries.size());
    for (const auto &e : entries) {
        w.put_string(e.key);
        w.put<uint64_t>(e.size_bytes);
    }
}
bool S2CSharedStateResponseP2P::read(Reader &r) {
    status = static_cast<SharedStateStatus>(r.get<uint8_t>());
    revision = r.get<uint64_t>();
    p2p = r.get<bool>();
    distributor = get_uuid(r);
    chunk_bytes = r.get<uint64_t>();
    const uint64_t n = r.get<uint64_t>();
    if (!r.plausible_count(n, 16)) return false;
    entries.clear();
    for (uint64_t i = 0; i < n && r.ok(); ++i) {
        SharedStateSlice s;
        s.key = r.get_string();
        s.size_bytes = r.get<uint64_t>();
        entries.push_back(std::move(s));
    }
    return r.ok();
}

// ---- benchmark
void C2BHello::write(Writer &w) const { put_uuid(w, uuid); }
bool C2BHello::read(Reader &r) {
    uuid = get_uuid(r);
    return r.ok();
}
void B2CBusy::write(Writer &w) const { w.put<bool>(busy); }
bool B2CBusy::read(Reader &r) {
    busy = r.get<bool>();
    return r.ok();
}

} // namespace pccl::wire
