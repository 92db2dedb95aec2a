/*
 * pccl.h -- public C99 interface of the B200-native PCCL build.
 *
 * The types, enumerator values and function signatures in the "core" part of this header are
 * ABI-compatible with the reference library (reference: include/pccl.h:22-458 and
 * ccoip/public_include/ccoip_inet.h:13-35), so a program written against the reference can be
 * recompiled against this header unchanged.  Differences in behaviour:
 *
 *   - send/recv buffers and shared-state tensors may be CUDA device pointers.  Peers that share
 *     one NVLink/NVSwitch domain then reduce with sm_100a kernels that read and write each
 *     other's memory directly; everything else falls back to the CCoIP TCP ring.
 *   - pcclFloat16 / pcclBFloat16 / pcclInt16 have their real element size (2 bytes) in
 *     shared-state bookkeeping (the reference computes 0, reference: src/pccl.cpp:525-543).
 *   - additional quantisation algorithms and a few extension entry points, all declared in the
 *     "extensions" part at the bottom and all prefixed so they cannot collide.
 */
#ifndef PCCL_H_
#define PCCL_H_

#ifdef __cplusplus
#include <cstddef>
#include <cstdint>
#else
#include <stdbool.h>
#include <stddef.h>
#include <stdint.h>
#endif

#if defined(_MSC_VER)
#define PCCL_EXPORT __declspec(dllexport)
#else
#define PCCL_EXPORT __attribute__((visibility("default")))
#endif

/* Marks pointer parameters that may be NULL. */
#define PCCL_NULLABLE

#ifdef __cplusplus
extern "C" {
#endif

/* ------------------------------------------------------------------------------------------- */
/* Network address types (layout identical to reference ccoip_inet.h)                           */
/* ------------------------------------------------------------------------------------------- */

typedef enum ccoip_inet_protocol_t { inetIPv4, inetIPv6 } ccoip_inet_protocol_t;

typedef struct ccoip_ipv4_address_t {
    uint8_t data[4];
} ccoip_ipv4_address_t;

typedef struct ccoip_ipv6_address_t {
    uint8_t data[16];
} ccoip_ipv6_address_t;

typedef struct ccoip_inet_address_t {
    ccoip_inet_protocol_t protocol;
    ccoip_ipv4_address_t ipv4;
    ccoip_ipv6_address_t ipv6;
} ccoip_inet_address_t;

typedef struct ccoip_socket_address_t {
    ccoip_inet_address_t inet;
    uint16_t port;
} ccoip_socket_address_t;

/* Default ports (reference: ccoip/public_include/ccoip.h:11-29). */
#define CCOIP_PROTOCOL_PORT_MASTER 48148
#define CCOIP_PROTOCOL_PORT_SHARED_STATE 48149
#define CCOIP_PROTOCOL_PORT_P2P 48150
#define CCOIP_PROTOCOL_PORT_BANDWIDTH_BENCHMARK 48151

/* ------------------------------------------------------------------------------------------- */
/* Core enums                                                                                   */
/* ------------------------------------------------------------------------------------------- */

typedef enum pcclResult_t {
    pcclSuccess = 0,
    pcclNotInitialized = 1,
    pcclInternalError = 2,
    pcclInvalidArgument = 3,
    pcclInvalidUsage = 4,
    pcclTooFewPeers = 5,
    pcclMasterConnectionFailed = 6,
    pcclRankConnectionFailed = 7,
    pcclRankConnectionLost = 8,
    pcclNoSharedStateAvailable = 9,
    pcclPendingAsyncOps = 10,
    pcclUpdateTopologyFailed = 11,
    pcclTopologyOptimizationFailed = 12
} pcclResult_t;

typedef enum pcclDataType_t {
    pcclUint8 = 0,
    pcclInt8 = 1,
    pcclInt16 = 2,
    pcclUint16 = 3,
    pcclUint32 = 4,
    pcclInt32 = 5,
    pcclUint64 = 6,
    pcclInt64 = 7,
    pcclFloat16 = 8,
    pcclBFloat16 = 9,
    pcclFloat = 10,
    pcclDouble = 11,
    /* extension: only legal as quantized_datatype together with pcclQuantBlockFp8 */
    pcclFloat8E4M3 = 12
} pcclDataType_t;

typedef enum pcclDeviceType_t {
    pcclDeviceCpu = 0,
    pcclDeviceCuda = 1
} pcclDeviceType_t;

typedef enum pcclRedOp_t { pcclSum, pcclAvg, pcclProd, pcclMax, pcclMin } pcclRedOp_t;

typedef enum pcclAttribute_t {
    /* number of peers in the whole run, over all peer groups */
    PCCL_ATTRIBUTE_GLOBAL_WORLD_SIZE = 1,
    /* number of peers in the caller's peer group */
    PCCL_ATTRIBUTE_PEER_GROUP_WORLD_SIZE = 2,
    /* number of different peer groups in the run */
    PCCL_ATTRIBUTE_NUM_DISTINCT_PEER_GROUPS = 3,
    /* size of the largest peer group */
    PCCL_ATTRIBUTE_LARGEST_PEER_GROUP_WORLD_SIZE = 4
} pcclAttribute_t;

/*
 * How a peer takes part in a shared-state synchronisation.
 *   ENFORCE_POPULAR  send or receive, whichever makes every peer hold the most popular content;
 *                    if any peer of a group uses it, all peers of that group must.
 *   RECEIVE_ONLY     never transmit; the caller's content does not take part in the election.
 *   SEND_ONLY        never be overwritten; the caller's content must be the elected one, peers
 *                    that declare SEND_ONLY with diverging content are kicked by the master.
 */
typedef enum pcclSharedStateSyncStrategy_t {
    PCCL_SHARED_STATE_SYNC_STRATEGY_ENFORCE_POPULAR = 0,
    PCCL_SHARED_STATE_SYNC_STRATEGY_RECEIVE_ONLY = 1,
    PCCL_SHARED_STATE_SYNC_STRATEGY_SEND_ONLY = 2
} pcclSharedStateSyncStrategy_t;

typedef enum pcclDistributionHint_t {
    pcclDistributionNone = 0,
    pcclDistributionNormal = 1,
    pcclDistributionUniform = 2
} pcclDistributionHint_t;

typedef enum pcclQuantizationAlgorithm_t {
    pcclQuantNone = 0,
    /* one (min,max) pair per transferred chunk, unsigned integer code words */
    pcclQuantMinMax = 1,
    /* one (zero point, scale) pair per transferred chunk, round-to-nearest */
    pcclQuantZeroPointScale = 2,
    /* extension: (min,max) per 128-element block, uint8 code words, fp32 accumulation */
    pcclQuantBlockMinMax = 3,
    /* extension: fp8-e4m3 code words with one fp32 scale per 128-element block */
    pcclQuantBlockFp8 = 4
} pcclQuantizationAlgorithm_t;

/* ------------------------------------------------------------------------------------------- */
/* Core structs                                                                                 */
/* ------------------------------------------------------------------------------------------- */

typedef struct {
    ccoip_socket_address_t master_address;
    uint32_t peer_group;
    uint32_t p2p_connection_pool_size;

    /* When true the three advertised_* addresses are handed to the master verbatim (NAT, port
     * forwarding); otherwise the master combines the source IP it sees with the ports below. */
    bool use_explicit_p2p_addresses;
    ccoip_socket_address_t advertised_p2p_address;
    ccoip_socket_address_t advertised_shared_state_address;
    ccoip_socket_address_t advertised_benchmark_address;

    /* Ports this peer binds locally.  A busy port is bumped upwards until one is free. */
    uint16_t internal_p2p_listen_port;          /* default 48149 */
    uint16_t internal_shared_state_listen_port; /* default 48150 */
    uint16_t internal_benchmark_listen_port;    /* default 48151 */
} pcclCommCreateParams_t;

typedef struct pcclComm_t pcclComm_t;
typedef struct pcclRankInfo_t pcclRankInfo_t;

typedef struct pcclReduceInfo_t {
    /* number of peers that took part in the (successful) operation */
    uint32_t local_world_size;
    uint64_t tx_bytes;
    uint64_t rx_bytes;
} pcclReduceInfo_t;

typedef struct pcclReduceOperandDescriptor_t {
    pcclDataType_t datatype;
    pcclDistributionHint_t distribution_hint;
} pcclReduceOperandDescriptor_t;

typedef struct pcclQuantizationOptions_t {
    pcclDataType_t quantized_datatype;
    pcclQuantizationAlgorithm_t algorithm;
} pcclQuantizationOptions_t;

typedef struct pcclReduceDescriptor_t {
    size_t count;
    pcclRedOp_t op;
    uint64_t tag;
    pcclReduceOperandDescriptor_t src_descriptor;
    pcclQuantizationOptions_t quantization_options;
} pcclReduceDescriptor_t;

typedef struct pcclReduceSingleDescriptor_t {
    void *sendbuf;
    void *recvbuf;
    pcclReduceDescriptor_t descriptor;
} pcclReduceOpDescriptor_t;

typedef struct pcclAsyncReduceOp_t {
    pcclComm_t *comm;
    uint64_t tag;
} pcclAsyncReduceOp_t;

/* Identity of a rank; never reused within a run. */
typedef struct pcclRankUuid_t {
    uint8_t data[16];
} pcclRankUuid_t;

typedef struct pcclTensorInfo_t {
    const char *name;
    void *data;
    size_t count;
    pcclDataType_t datatype;
    pcclDeviceType_t device_type;
    bool allow_content_inequality;
} pcclTensorInfo_t;

typedef struct pcclSharedState_t {
    uint64_t revision;
    size_t count;
    pcclTensorInfo_t *infos;
} pcclSharedState_t;

typedef struct pcclSharedStateSyncInfo_t {
    uint64_t tx_bytes;
    uint64_t rx_bytes;
} pcclSharedStateSyncInfo_t;

typedef struct pcclMasterInstanceState_t pcclMasterInstance_t;

typedef struct pcclBuildInfo_t {
    bool has_cuda_support;
} pcclBuildInfo_t;

/* ------------------------------------------------------------------------------------------- */
/* Core functions (reference: include/pccl.h:225-458)                                           */
/* ------------------------------------------------------------------------------------------- */

/* Must be the first call into the library. */
PCCL_EXPORT pcclResult_t pcclInit(void);

/* Allocates a communicator; no network activity happens before pcclConnect. */
PCCL_EXPORT pcclResult_t pcclCreateCommunicator(const pcclCommCreateParams_t *params, pcclComm_t **comm_out);

/* World-size style attributes; values are those of the last completed topology update. */
PCCL_EXPORT pcclResult_t pcclGetAttribute(const pcclComm_t *communicator, pcclAttribute_t attribute,
                                          int *p_attribute_out);

/* Closes all sockets, joins all threads, releases device resources. */
PCCL_EXPORT pcclResult_t pcclDestroyCommunicator(pcclComm_t *communicator);

/* Joins the run.  Blocks until the already accepted peers vote to admit this peer. */
PCCL_EXPORT pcclResult_t pcclConnect(pcclComm_t *communicator);

/* Joint vote of all accepted peers to admit pending peers and (re)wire P2P links.
 * Fails with pcclPendingAsyncOps while async reductions are in flight. */
PCCL_EXPORT pcclResult_t pcclUpdateTopology(pcclComm_t *communicator);

/* Joint query whether peers wait for admission; legal while reductions are in flight. */
PCCL_EXPORT pcclResult_t pcclArePeersPending(const pcclComm_t *communicator, bool *pending_out);

/* Joint bandwidth measurement + ring re-ordering (ATSP over measured link speeds). */
PCCL_EXPORT pcclResult_t pcclOptimizeTopology(const pcclComm_t *communicator);

/* Blocking all-reduce.  sendbuff may equal recvbuff.  On failure the contents of sendbuff are
 * what they were before the call. */
PCCL_EXPORT pcclResult_t pcclAllReduce(const void *sendbuff, void *recvbuff, const pcclReduceDescriptor_t *descriptor,
                                       const pcclComm_t *communicator, pcclReduceInfo_t *PCCL_NULLABLE reduce_info_out);

/* Starts an all-reduce; completion is observed with pcclAwaitAsyncReduce.  A tag may only be in
 * flight once. */
PCCL_EXPORT pcclResult_t pcclAllReduceAsync(const void *sendbuff, void *recvbuff,
                                            const pcclReduceDescriptor_t *descriptor, const pcclComm_t *communicator,
                                            pcclAsyncReduceOp_t *reduce_handle_out);

/* Runs a list of reductions with at most max_in_flight concurrently; failed ones are retried
 * with the surviving peers until all succeeded or fewer than two peers are left. */
PCCL_EXPORT pcclResult_t pcclAllReduceMultipleWithRetry(const pcclReduceOpDescriptor_t *descriptors, size_t count,
                                                        const pcclComm_t *communicator,
                                                        pcclReduceInfo_t *PCCL_NULLABLE reduce_info_out,
                                                        int max_in_flight);

PCCL_EXPORT pcclResult_t pcclAwaitAsyncReduce(const pcclAsyncReduceOp_t *reduce_handle,
                                              pcclReduceInfo_t *PCCL_NULLABLE reduce_info_out);

/* Brings every peer of the group to the elected ("most popular") shared-state content; only
 * tensors whose hash differs are transferred.  shared_state->revision may be updated. */
PCCL_EXPORT pcclResult_t pcclSynchronizeSharedState(const pcclComm_t *communicator, pcclSharedState_t *shared_state,
                                                    pcclSharedStateSyncStrategy_t strategy,
                                                    pcclSharedStateSyncInfo_t *PCCL_NULLABLE sync_info_out);

PCCL_EXPORT pcclResult_t pcclCreateMaster(ccoip_socket_address_t listen_address,
                                          pcclMasterInstance_t **p_master_handle_out);
/* Non-blocking: the master runs on its own thread. */
PCCL_EXPORT pcclResult_t pcclRunMaster(pcclMasterInstance_t *master_instance);
PCCL_EXPORT pcclResult_t pcclInterruptMaster(pcclMasterInstance_t *master_instance);
PCCL_EXPORT pcclResult_t pcclMasterAwaitTermination(pcclMasterInstance_t *master_instance);
PCCL_EXPORT pcclResult_t pcclDestroyMaster(pcclMasterInstance_t *master_instance);

PCCL_EXPORT pcclResult_t pcclGetBuildInfo(pcclBuildInfo_t *info);

/* ------------------------------------------------------------------------------------------- */
/* Extensions (no counterpart in the reference)                                                 */
/* ------------------------------------------------------------------------------------------- */

/* Which data plane an all-reduce used. */
typedef enum pcclExtDataPlane_t {
    pcclExtPlaneNone = 0,
    pcclExtPlaneTcpRing = 1,     /* CCoIP ring over sockets (host memory, staged for device buffers) */
    pcclExtPlaneP2POneShot = 2,  /* every GPU reads every peer, latency regime */
    pcclExtPlaneP2PTwoShot = 3,  /* reduce-scatter + all-gather with direct peer loads */
    pcclExtPlaneNvls = 4,        /* NVSwitch multicast ld_reduce / st */
    pcclExtPlaneTcpHierarchical = 5 /* two-level ring for groups that span hosts: local reduce-scatter, cross-host ring, local all-gather */
} pcclExtDataPlane_t;

typedef struct pcclExtReduceStats_t {
    pcclExtDataPlane_t plane;
    float device_ms;     /* CUDA-event time of the device part, 0 for host ops */
    uint64_t seq_nr;     /* master-issued sequence number of the operation */
    uint32_t kernel_launches; /* kernels of this library launched for the operation */
    uint32_t leased;     /* 1: ran under a lease, without a vote at the master (seq_nr then counts leased operations of the tag) */
    uint64_t phase_ns[8]; /* device plane: GPU global timer (ns) when block 0 of the (last) kernel reached its phase boundaries:
                             [0] start, [1] input staged, [2] first barrier passed, [3] own chunk reduced, [4] second barrier
                             passed, [5] output written, [6] closing barrier passed; 0 = phase not present */
    uint32_t has_content_hash; /* 1: the kernel produced the mset64 content hash of what it wrote as a by-product */
    uint64_t content_hash[2];  /* all-reduce: [0] hash of the result; outer step: [0] new parameters, [1] new momentum */
} pcclExtReduceStats_t;

/* Stats of the most recently completed reduction with the given tag. */
PCCL_EXPORT pcclResult_t pcclExtGetLastReduceStats(const pcclComm_t *communicator, uint64_t tag,
                                                   pcclExtReduceStats_t *stats_out);

/* Port the master actually bound (useful with port 0 / bumped ports in tests). */
PCCL_EXPORT pcclResult_t pcclExtMasterGetPort(const pcclMasterInstance_t *master_instance, uint16_t *port_out);

/* Hash of a host or device buffer with the hash the shared-state protocol uses.
 * hash_type: 0 = simplehash (reference-compatible value), 1 = CRC32C, 2 = mset64 (order-independent sum of mixed
 * 16-byte packets; the one the collective kernels can produce as a by-product). */
PCCL_EXPORT pcclResult_t pcclExtHashBuffer(const void *data, size_t n_bytes, pcclDeviceType_t device_type,
                                           int hash_type, uint64_t *hash_out);

/* Allocates `bytes` of device memory on the current CUDA device from the communicator's
 * peer-visible arena.  Reductions whose send/recv buffers both come from here skip the
 * staging copy.  Freed by pcclExtSymmetricFree or with the communicator. */
PCCL_EXPORT pcclResult_t pcclExtSymmetricAlloc(pcclComm_t *communicator, size_t bytes, void **ptr_out);
PCCL_EXPORT pcclResult_t pcclExtSymmetricFree(pcclComm_t *communicator, void *ptr);

/* Size of one element of a data type in bytes, 0 for unknown values. */
PCCL_EXPORT size_t pcclExtDataTypeSize(pcclDataType_t datatype);

/* Stream-ordered all-reduce on device buffers (no counterpart in the reference, whose all-reduce takes host memory and
 * blocks: src/pccl.cpp:301-313).  The operation is ordered after everything already queued on `cuda_stream`
 * (a cudaStream_t; NULL = the legacy default stream) and that stream waits for the result, so the caller needs no
 * stream or device synchronisation on either side and kernels on other streams keep running while peers exchange data.
 * enqueue != 0: when the operation needs no vote at the master (it runs under a lease) the call
 * returns right after the launch; its verdict is taken by the next stream-ordered operation on the same tag or by
 * pcclExtWaitStreamOps, and until then the tag counts as in flight (pcclUpdateTopology and
 * pcclSynchronizeSharedState answer pcclPendingAsyncOps).  A failure of an enqueued operation is reported by that
 * later call as pcclRankConnectionLost. */
PCCL_EXPORT pcclResult_t pcclExtAllReduceOnStream(const void *sendbuff, void *recvbuff, const pcclReduceDescriptor_t *descriptor,
                                                  const pcclComm_t *communicator, void *cuda_stream, int enqueue,
                                                  pcclReduceInfo_t *reduce_info_out);
/* Takes the verdict of the enqueued operation on `tag` (of all tags when all_tags != 0); blocks until it has finished. */
PCCL_EXPORT pcclResult_t pcclExtWaitStreamOps(const pcclComm_t *communicator, uint64_t tag, int all_tags,
                                              pcclReduceInfo_t *reduce_info_out);

/* Content hashes as a by-product (no counterpart in the reference, which hashes every shared-state tensor in a pass of its own
 * before each synchronisation: ccoip_client_handler.cpp:383-416).  With trust switched on, the fused DiLoCo outer step hashes
 * the parameters and the momentum buffer while it writes them (hash type mset64, PCCL_SHARED_STATE_HASH=mset), and the next
 * pcclSynchronizeSharedState takes the hash of a tensor this library wrote last from that record instead of launching a hash
 * kernel.  The caller promises not to modify such tensors behind the library's back; anything the library itself writes
 * (reductions, received shared state) drops the record.  pcclExtHashKernelsSkipped counts the hashes that were saved. */
PCCL_EXPORT pcclResult_t pcclExtTrustContentHashes(pcclComm_t *communicator, int on);
PCCL_EXPORT pcclResult_t pcclExtHashKernelsSkipped(const pcclComm_t *communicator, uint64_t *count_out);

/* DiLoCo outer step as one collective operation (no counterpart in the reference, whose example performs the
 * three stages with torch ops on the host around pcclAllReduce: python/examples/nanogpt_diloco/sync_diloco.py:480-530):
 *     g        = average over all peers of (params - inner)
 *     momentum = first_step ? g : mu * momentum + g
 *     params  -= lr * (g + mu * momentum)            (torch.optim.SGD with nesterov=True)
 *     inner    = params
 * All three tensors are float32[count], all on the host or all on one CUDA device.  Between GPUs of one NVLink
 * domain the whole step is a single kernel per peer (pseudo-gradient while staging, reduction over NVLink /
 * NVSwitch, update while writing out); otherwise -- TCP ring, or more elements than one staging slot holds --
 * the library computes the pseudo-gradient into a scratch buffer, reduces that, and applies the update.
 * Either way the tensors are untouched when the call fails (pcclRankConnectionLost: retry after
 * pcclUpdateTopology, like pcclAllReduce).  Synchronous; one tag like any reduction. */
typedef struct pcclExtOuterStepDescriptor_t {
    void *params;
    void *inner;
    void *momentum;
    uint64_t count;
    float lr;
    float mu;
    int first_step; /* non-zero: momentum holds nothing yet */
    uint64_t tag;
} pcclExtOuterStepDescriptor_t;
PCCL_EXPORT pcclResult_t pcclExtDiLoCoOuterStep(const pcclExtOuterStepDescriptor_t *descriptor, const pcclComm_t *communicator,
                                                pcclReduceInfo_t *PCCL_NULLABLE reduce_info_out);

#ifdef __cplusplus
}
#endif

#endif /* PCCL_H_ */
