/*
 * ptype_b200.h — C ABI of the B200 batched-RPC-dispatch path for edegens/ptype.
 *
 * This is the whole drop-in boundary: plain C, plain pointers and sizes, `int`
 * status returns (0 = OK, never an exception), no torch / C++ types.  A Go host
 * binds it through cgo; this repo's tests and bench bind it
 * through Python ctypes.  Citations are file:line in the reference tree
 * (edegens/ptype) and name the code each item replaces or mirrors.
 *
 * One batch = N independent calls.  For each call the device does what one
 *   cluster.(*Client).Call               (cluster/rpc.go:59-67)
 * does through net/rpc today:
 *   1. find the service's node list in the registry snapshot
 *        etcdRegistry.Services / nodes   (cluster/registry.go:93-117,152-166)
 *   2. apply the node-selection rule
 *        selectNodes + hashConnectionIndex (cluster/rpc.go:252-270)
 *        roundRobinSelect                  (cluster/rpc.go:176-183)
 *   3. resolve "Receiver.Method" the way the net/rpc server does
 *        (Go 1.13 stdlib net/rpc; call sites example/calculator/server/server.go:16-20)
 *   4. run the method body
 *        Calculator.Multiply             (example/calculator/calculator.go:9-12)
 *        Calculator.Add                  (not in the reference)
 *        Prime.Check                     (example/optimus/prime.go:15-25)
 *   5. pack a fixed-size reply record.
 */
#ifndef PTYPE_B200_H
#define PTYPE_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PTYPE_B200_ABI_VERSION 1

/* ------------------------------------------------------------------ records */

/* One call attempt.  48 bytes, 16-byte aligned: three 128-bit loads.
 * Replaces the (serviceMethod, args) pair of Client.Call (cluster/rpc.go:59)
 * plus the balancer's sequence number (cluster/rpc.go:181). */
typedef struct ptype_req {
    uint64_t svc_key; /* ptype_b200_service_key(cluster service name): the name a
                         Client was created for (cluster/cluster.go:101-103)       */
    uint64_t seq;     /* value atomic.AddUint64(&c.seq,1) returned for this attempt
                         (cluster/rpc.go:181); a batch of N on one Client carries
                         base+1 .. base+N                                           */
    uint32_t method;  /* FNV-1a-32 of the whole "Receiver.Method" string           */
    uint32_t recv;    /* FNV-1a-32 of the text before the last '.', or
                         PTYPE_RECV_ILL_FORMED when the string has no '.'          */
    int64_t a0, a1, a2; /* calculator.Args{A,B} (example/calculator/calculator.go:3-5)
                           or prime.Args{Min,Max,Target} (example/optimus/prime.go:7-11) */
} ptype_req;

/* One reply.  16 bytes: one 128-bit store; reply j belongs to request j. */
typedef struct ptype_reply {
    int64_t value;   /* *reply; 0 when status != OK (the caller's reply is left
                        untouched on error, cluster/rpc_test.go:160-186)          */
    uint32_t node;   /* routing decision: global node id in the snapshot (rank of the
                        node's key in the sorted `services` range), or PTYPE_NO_NODE */
    uint32_t status; /* enum ptype_status                                          */
} ptype_reply;

#define PTYPE_NO_NODE 0xFFFFFFFFu
#define PTYPE_RECV_ILL_FORMED 0xFFFFFFFFu

/* Per-call status.  What the Go caller would have seen from Client.Call. */
enum ptype_status {
    PTYPE_ST_OK = 0,
    PTYPE_ST_NO_SERVICE = 1,        /* service absent from the snapshot: its watch delivers an
                                       empty list, Call returns ErrNoClientAvailable
                                       (cluster/rpc.go:16,62-64,178-180)                        */
    PTYPE_ST_NO_CLIENT = 2,         /* service present with an empty list (same Go error)       */
    PTYPE_ST_ILL_FORMED = 3,        /* "rpc: service/method request ill-formed: X" [net/rpc]    */
    PTYPE_ST_CANT_FIND_SERVICE = 4, /* "rpc: can't find service X" [net/rpc]                    */
    PTYPE_ST_CANT_FIND_METHOD = 5,  /* "rpc: can't find method X" [net/rpc]                     */
    PTYPE_ST_BAD_INDEX = 6,         /* seq >= 2^63 and int(seq)%k != 0: clients[negative] panics
                                       in Go (cluster/rpc.go:182)                               */
    PTYPE_ST_RANGE_LIMIT = 7,       /* Prime.Check range wider than the context's iteration
                                       limit; not executed (no counterpart: the reference
                                       would loop, sleeping 250 ms per step, prime.go:16-17)   */
    PTYPE_ST__COUNT = 8
};

/* Built-in method bodies (opcodes in the image's method table). */
enum ptype_opcode {
    PTYPE_OP_NONE = 0,
    PTYPE_OP_MULTIPLY = 1, /* example/calculator/calculator.go:9-12 */
    PTYPE_OP_ADD = 2,      /* SURVEY.md D1 */
    PTYPE_OP_CHECK = 3     /* example/optimus/prime.go:15-25 (sleep dropped) */
};

/* ------------------------------------------------------- registry image blob */
/* Little-endian, position independent, self-describing; the same bytes go to
 * every GPU (H2D or ncclBroadcast).  It is the device form of what
 * Registry.Services() returns (cluster/registry.go:93-117) plus the inputs of
 * the selection rule (localAddr, ConnConfig.MaxConnections; cluster/rpc.go:19-38,128). */

#define PTYPE_IMAGE_MAGIC 0x49595450u /* "PTYI" */
#define PTYPE_IMAGE_VERSION 4u
#define PTYPE_SLOT_WAYS 3        /* candidate slots per key (3-way cuckoo hashing) */
#define PTYPE_MTABLE_SLOTS 64   /* perfect-hash method table size                  */
#define PTYPE_SLOT_EMPTY 0xFFFFFFFFu
#define PTYPE_MAX_RECEIVERS 8
#define PTYPE_MAX_METHODS 32
#define PTYPE_LOCAL_ADDR_MAX 64

typedef struct ptype_image_header { /* 192 bytes */
    uint32_t magic, version;
    uint64_t epoch;          /* snapshot generation, chosen by the publisher          */
    uint64_t total_bytes;    /* whole blob                                             */
    uint64_t body_checksum;  /* 4-lane word-wise FNV of blob[192 .. total_bytes)        */
    uint64_t key_salt;       /* xor-ed into the FNV-1a-64 offset basis (0 = plain FNV) */
    uint32_t n_services;     /* occupied slots                                         */
    uint32_t n_slots;        /* cuckoo table size (any value >= 1, no power of two)    */
    uint32_t n_nodes;        /* global node array length                               */
    uint32_t max_connections;/* ConnConfig.MaxConnections; 0 = mesh (rpc.go:20-22)     */
    uint32_t n_conn_hash;    /* == max_connections                                     */
    uint32_t n_methods;      /* <= PTYPE_MAX_METHODS                                   */
    uint32_t n_receivers;    /* <= PTYPE_MAX_RECEIVERS                                 */
    uint32_t off_slots;      /* ptype_slot[n_slots], 16-byte aligned                   */
    uint32_t off_conn_hash;  /* uint32[n_conn_hash]: FNV-1a-32(localAddr+itoa(i))      */
    uint32_t off_methods;    /* ptype_method[n_methods]                                */
    uint32_t off_nodes;      /* ptype_node[n_nodes]      (host-side lookups only)      */
    uint32_t off_services;   /* ptype_service[n_services](host-side lookups only)      */
    uint32_t off_strings;    /* string pool, NUL-terminated entries                    */
    uint32_t strings_bytes;
    uint32_t device_bytes;   /* prefix of the blob the kernels read (slots, hashes, methods, order) */
    uint32_t table_seed;     /* xor-ed into the folded key before the three slot hashes */
    uint32_t method_mul;     /* odd multiplier: mtable index = (method_hash * mul) >> 26 */
    uint32_t off_mtable;     /* ptype_method[PTYPE_MTABLE_SLOTS], collision-free, opcode 0 = empty */
    uint32_t host_sections;  /* 1: services/nodes/strings present; 0: device-only image       */
    uint32_t off_order;      /* uint32[n_slots]: rank of the slot's service in etcd key order
                                (services/<name>/ compared bytewise), PTYPE_SLOT_EMPTY for a free
                                slot; read by ptype_b200_update_service only                       */
    uint32_t reserved[2];
    char local_addr[PTYPE_LOCAL_ADDR_MAX]; /* the Client host's address string (cluster.go:69,82) */
} ptype_image_header;

/* One table slot, 16 bytes (one LDS.128 / LDG.128).  A key lives in exactly one of its
 * PTYPE_SLOT_WAYS candidate slots (ptype::slot_of in csrc/table_hash.h), so a lookup is
 * at most three probes, hit or miss. */
typedef struct ptype_slot {
    uint64_t key;       /* ptype_b200_service_key(name, key_salt)                      */
    uint32_t node_base; /* global id of the service's first node; PTYPE_SLOT_EMPTY = free */
    uint32_t n_and_recv;/* bits 0..23: number of registered nodes n (sorted list length)
                           bits 24..31: bitmask of receivers its nodes registered      */
} ptype_slot;

typedef struct ptype_method { /* 16 bytes */
    uint32_t method_hash; /* FNV-1a-32("Receiver.Method") */
    uint32_t recv_hash;   /* FNV-1a-32("Receiver")        */
    uint32_t opcode;      /* enum ptype_opcode; PTYPE_OP_NONE = receiver-only entry */
    uint32_t recv_bit;    /* bit index in ptype_slot.n_and_recv >> 24               */
} ptype_method;

typedef struct ptype_node { /* host-side: cluster.Node (cluster/registry.go:23-26) */
    uint32_t address_off; /* into the string pool */
    uint32_t name_off;    /* node name (last key element) */
    uint32_t service;     /* index into ptype_service[] */
    int32_t port;
} ptype_node;

typedef struct ptype_service { /* host-side, sorted by key bytes like the etcd range */
    uint64_t key;
    uint32_t name_off;
    uint32_t node_base;
    uint32_t n_nodes;
    uint32_t recv_mask;
} ptype_service;

/* ------------------------------------------------------------------- errors */
enum ptype_err {
    PTYPE_OK = 0,
    PTYPE_E_INVAL = 1,     /* bad argument                                   */
    PTYPE_E_NOMEM = 2,
    PTYPE_E_CUDA = 3,      /* a CUDA call failed; text via ptype_b200_last_error */
    PTYPE_E_NO_IMAGE = 4,  /* dispatch before any publish_image              */
    PTYPE_E_BAD_IMAGE = 5, /* magic / version / checksum / bounds            */
    PTYPE_E_COLLISION = 6, /* builder could not find a collision-free salt   */
    PTYPE_E_NCCL = 7,
    PTYPE_E_LIMIT = 8,     /* size beyond a documented maximum               */
    PTYPE_E_NOT_FOUND = 9,
    PTYPE_E_BUSY = 10
};
const char* ptype_b200_strerror(int err);
/* Text of the last CUDA/NCCL failure on the calling thread ("" if none). */
const char* ptype_b200_last_error(void);
int ptype_b200_abi_version(void);

/* --------------------------------------------------------- hashing (host) */
/* hash/fnv New32a / New64a [Go stdlib]; the only explicit hash in the reference
 * is fnv.New32a at cluster/rpc.go:267. */
uint32_t ptype_b200_fnv1a32(const void* data, size_t len);
uint64_t ptype_b200_fnv1a64(const void* data, size_t len);
/* Key of a cluster service name (the string passed to Cluster.NewClient,
 * cluster/cluster.go:101).  salt comes from the image header (0 unless two
 * registered names collide). */
uint64_t ptype_b200_service_key(const char* service_name, uint64_t salt);
/* Split "Receiver.Method" at the last '.' like net/rpc's server does and hash
 * both parts.  No '.' → *recv = PTYPE_RECV_ILL_FORMED. */
void ptype_b200_method_key(const char* service_method, uint32_t* method, uint32_t* recv);
/* hashConnectionIndex (cluster/rpc.go:266-270). */
uint32_t ptype_b200_conn_hash(const char* local_addr, int conn_number);

/* --------------------------------------------------- image builder (host) */
/* Collects what Registry.Register wrote (cluster/registry.go:51-86) and emits
 * the blob.  Needs no GPU. */
typedef struct ptype_builder ptype_builder;
/* local_addr: the Client host's address (cluster/cluster.go:69,82; tests use "").
 * max_connections: ConnConfig.MaxConnections (3 by default, 0 = mesh; <0 rejected). */
int ptype_b200_builder_new(const char* local_addr, int max_connections, ptype_builder** out);
/* One registered node = one etcd key services/<service>/<node_name>/ with value
 * {"address":address,"port":port}.  Re-registering a key overwrites it. */
int ptype_b200_builder_register(ptype_builder*, const char* service, const char* node_name,
                                const char* address, int port);
/* The key of one node went away (lease expiry, cluster/registry.go:58-83, or a delete as in
 * cluster/registry_test.go:179-191).  PTYPE_E_NOT_FOUND if it was not registered. */
int ptype_b200_builder_unregister(ptype_builder*, const char* service, const char* node_name);
/* Emit only what the kernels read (slots, connection hashes, method tables): small blobs for
 * frequent live updates.  ptype_b200_image_selected / _node answer PTYPE_E_NOT_FOUND on them. */
int ptype_b200_builder_set_device_only(ptype_builder*, int on);
/* A service whose watch delivered an empty list (cluster/rpc_test.go:85-90). */
int ptype_b200_builder_declare(ptype_builder*, const char* service);
/* The nodes of `service` ran rpc.Register(new(<receiver>)) (server.go:16-17).
 * Known receivers: "Calculator" {Multiply, Add}, "Prime" {Check}. */
int ptype_b200_builder_expose(ptype_builder*, const char* service, const char* receiver);
/* Override the open-addressing table size (must exceed the number of services).
 * Default: half full, or sized to fill the kernel's shared-memory budget. */
int ptype_b200_builder_set_slots(ptype_builder*, uint32_t n_slots);
/* Emit the blob; caller frees with ptype_b200_image_free. */
int ptype_b200_builder_finish(ptype_builder*, uint64_t epoch, void** image, size_t* bytes);
void ptype_b200_builder_free(ptype_builder*);
void ptype_b200_image_free(void* image);
int ptype_b200_image_validate(const void* image, size_t bytes);
/* Host-side view of a snapshot: what getSelectedNodes() (cluster/rpc.go:293-297)
 * would hold for a Client of `service`.  selected[] receives up to cap global
 * node ids; *k the list length. */
int ptype_b200_image_selected(const void* image, const char* service, uint32_t* selected,
                              size_t cap, uint32_t* k);
/* cluster.Node of a global node id. */
int ptype_b200_image_node(const void* image, uint32_t node_id, char* address, size_t address_cap,
                          int* port, char* service, size_t service_cap, char* node_name,
                          size_t node_name_cap);

/* ------------------------------------------------------------------ context */
typedef struct ptype_ctx ptype_ctx;
/* Creates one stream set and a double-buffered image slot per listed CUDA
 * device.  Entry points may be called from any OS thread (cgo): device state is
 * held in the context, never in thread-locals. */
int ptype_b200_open(const int* devices, int ndev, ptype_ctx** out);
void ptype_b200_close(ptype_ctx*);
int ptype_b200_device_count(int* out);

/* Pinned host staging the caller may fill (Go slices must not be retained
 * across an async call; this memory can). */
void* ptype_b200_alloc_pinned(size_t bytes);
void ptype_b200_free_pinned(void*);
/* Device buffers for callers that keep batches resident in HBM. */
int ptype_b200_device_alloc(ptype_ctx*, int dev, size_t bytes, void** out);
int ptype_b200_device_free(ptype_ctx*, int dev, void* p);
int ptype_b200_memcpy_h2d(ptype_ctx*, int dev, void* dst, const void* src, size_t bytes);
int ptype_b200_memcpy_d2h(ptype_ctx*, int dev, void* dst, const void* src, size_t bytes);
/* The cudaStream_t dispatches of device `dev` are ordered on (for event timing, or to enqueue a
 * kernel of your own that fills a batch).  Ordering rule: consecutive dispatch launches of this
 * library overlap each other's ramp-up and drain (programmatic dependent launch), which is sound
 * because a dispatch never reads what the previous one wrote.  The library drops that overlap by
 * itself after anything else IT enqueues on the stream (copies, assign_seq, codec and checksum
 * kernels, a new image).  Work YOU enqueue there is invisible to it: if that work produces the
 * requests (or anything else) the next dispatch reads, pass PTYPE_F_AFTER_PRODUCER to that
 * dispatch so that it starts only after everything before it on the stream has completed. */
int ptype_b200_stream(ptype_ctx*, int dev, void** cuda_stream);

/* Install a new snapshot on every device of the context: copy into the inactive
 * image buffer, then flip at a batch boundary.  Mirrors the pointer swap of
 * handleNewNodes (cluster/rpc.go:233-236).  An image with zero services is
 * installed as given; "ignore empty updates" (rpc.go:204-206) is host policy. */
int ptype_b200_publish_image(ptype_ctx*, const void* image, size_t bytes);
/* One node joined (+1) or left (-1) `service` (any delta is accepted): patch the installed
 * snapshot ON THE DEVICE instead of rebuilding and uploading an image.  The service's node count
 * changes by delta and every service that sorts after it moves its node ids by delta - what
 * rebuilding from the changed registry would give (cluster/registry.go:152-166 re-lists the
 * service on every watch event; this is that event applied in place).  The active table is
 * copied to the inactive buffer and patched there, then the epoch advances by one at a batch
 * boundary, without stopping the dispatch streams.  In a multi-process deployment every
 * process applies the same events (each has its own watch, as in the reference).
 * PTYPE_E_NOT_FOUND: the snapshot has no such service (a new service needs a new image);
 * PTYPE_E_LIMIT: the count would leave [0, 2^24).  Host-side lookups on the context's copy of
 * the image (ptype_b200_current_image) see the patched table but lose the node names. */
int ptype_b200_update_service(ptype_ctx*, const char* service, int delta);
int ptype_b200_epoch(ptype_ctx*, int dev, uint64_t* epoch);
/* Copy of the blob last installed (on non-root ranks: the bytes NCCL delivered).
 * *bytes in: capacity of out, out: size of the blob; PTYPE_E_LIMIT if too small. */
int ptype_b200_current_image(ptype_ctx*, void* out, size_t* bytes);

/* Multi-process (one process per GPU): NCCL is used only here.  Rank `root`
 * passes the blob, the others pass NULL/0; every rank ends with the same
 * snapshot installed.  The 128-byte id is created on one rank and carried to
 * the others by the host (in ptype: a KVStore.Put, cluster/store.go:56). */
int ptype_b200_nccl_unique_id(void* id128);
int ptype_b200_comm_init(ptype_ctx*, const void* id128, int rank, int nranks);
int ptype_b200_broadcast_image(ptype_ctx*, const void* image, size_t bytes, int root);

/* ----------------------------------------------------------------- dispatch */
#define PTYPE_F_DEVICE_PTRS 1u /* reqs/replies are device pointers on `dev`      */
#define PTYPE_F_ASYNC 2u       /* return after enqueue; pair with ptype_b200_sync */
#define PTYPE_F_GLOBAL_TABLE 4u/* measurement switch: probe the table in L2, not shared memory  */
#define PTYPE_F_LDG_LOADS 8u   /* measurement switch: fetch requests with LDG + shuffle transpose */
#define PTYPE_F_BULK_LOADS 16u /* measurement switch: stage requests with bulk-async copies      */
#define PTYPE_F_AFTER_PRODUCER 64u /* the work enqueued before this call on ptype_b200_stream() wrote
                                     this batch's inputs: no overlap with it (see ptype_b200_stream) */
#define PTYPE_F_ONE_LAUNCH_PER_BATCH 32u /* dispatch_batches: one kernel launch per batch instead of
                                            one launch per group of up to 96 batches               */

/* Optional per-batch outputs (all device-written; host copies made on sync). */
typedef struct ptype_batch_stats {
    uint64_t calls;
    uint64_t by_status[PTYPE_ST__COUNT];
    uint64_t failed_count; /* entries written to the failed-call list */
} ptype_batch_stats;

/* N calls through the snapshot currently installed on `dev`.
 * Host pointers: pageable or pinned; copies are pipelined in chunks.
 * Device pointers (PTYPE_F_DEVICE_PTRS): no copies, one kernel launch.
 * Limit: Prime.Check ranges longer than the context's check limit get
 * PTYPE_ST_RANGE_LIMIT (ptype_b200_set_check_limit; default 1<<16). */
int ptype_b200_dispatch(ptype_ctx*, int dev, const ptype_req* reqs, size_t n, ptype_reply* replies,
                        uint32_t flags);
/* Same, and also compacts the indices of calls with status != OK into
 * failed_idx[0..failed_cap) (order unspecified) and fills *stats — the set a
 * withRetry loop (cluster/rpc.go:107-116) would re-issue.  failed_idx/stats
 * follow the same host/device convention as reqs. */
int ptype_b200_dispatch_ex(ptype_ctx*, int dev, const ptype_req* reqs, size_t n,
                           ptype_reply* replies, uint32_t* failed_idx, size_t failed_cap,
                           ptype_batch_stats* stats, uint32_t flags);
/* Several device-resident batches in one call.  Groups of up to 96 batches run as ONE
 * kernel launch whose persistent grid walks all their tiles as a single stream (no gap or
 * ramp between batches; for a list of 2-48 batches over a small table that finds the stream
 * idle, the last 8 % of the tiles follow as a second launch of short CTAs that evens out the end); with PTYPE_F_ONE_LAUNCH_PER_BATCH, or when the registry table is
 * probed in L2, every batch is its own launch and consecutive kernels overlap through
 * programmatic dependent launch.  Returns without waiting when PTYPE_F_ASYNC is set.  This is what a host pipelining the batches of many Clients
 * (one Client per service, cluster/cluster.go:101-103) submits per scheduling round. */
typedef struct ptype_batch {
    const ptype_req* reqs; /* device pointer */
    ptype_reply* replies;  /* device pointer */
    size_t n;
} ptype_batch;
int ptype_b200_dispatch_batches(ptype_ctx*, int dev, const ptype_batch* batches, size_t count,
                                uint32_t flags);
/* One Client's batch in its compact form — what (*Client).CallBatch would submit.  All n
 * calls name the service `svc_key` and own the consecutive sequence numbers
 * seq_base+1 .. seq_base+n (ONE atomic.AddUint64(&c.seq, n), cluster/rpc.go:181), so only the
 * arguments travel per call: args holds args_per_call (2: calculator.Args{A,B}; 3:
 * prime.Args{Min,Max,Target}) int64 per call.  method_sel is NULL when every call uses
 * methods[0], else one byte per call indexing methods[] (values >= n_methods read as 0).
 * Replies are the same 16-byte records as ptype_b200_dispatch and identical to what
 * dispatching the equivalent 48-byte ptype_req records gives.  Host pointers are copied in
 * pipelined chunks (16-25 B/call over PCIe instead of 48); PTYPE_F_DEVICE_PTRS = one launch. */
#define PTYPE_BATCH_METHODS 4
typedef struct ptype_method_ref {
    uint32_t method; /* as ptype_req.method */
    uint32_t recv;   /* as ptype_req.recv   */
} ptype_method_ref;
int ptype_b200_call_batch(ptype_ctx*, int dev, uint64_t svc_key, uint64_t seq_base,
                          const ptype_method_ref* methods, int n_methods, const uint8_t* method_sel,
                          const int64_t* args, int args_per_call, size_t n, ptype_reply* replies,
                          uint32_t flags);
/* The same batch with the replies CallBatch's caller actually reads: values[j] (8 bytes per
 * call, 0 for a failed call) and, for the calls whose status is not OK, one (index, status)
 * entry in failed[0..failed_cap) (order unspecified; *failed_count is the exact number even if
 * it exceeds failed_cap).  Routing decisions are not returned.  Host pointers (pinned: the
 * kernel streams across PCIe itself; pageable: staged copies) or, with PTYPE_F_DEVICE_PTRS,
 * device pointers for args/method_sel/values; failed and failed_count are always host. */
typedef struct ptype_failed {
    uint32_t index;
    uint32_t status; /* enum ptype_status */
} ptype_failed;
int ptype_b200_call_batch_values(ptype_ctx*, int dev, uint64_t svc_key, uint64_t seq_base,
                                 const ptype_method_ref* methods, int n_methods, const uint8_t* method_sel,
                                 const int64_t* args, int args_per_call, size_t n, int64_t* values,
                                 ptype_failed* failed, size_t failed_cap, uint64_t* failed_count,
                                 uint32_t flags);
int ptype_b200_sync(ptype_ctx*, int dev);
int ptype_b200_set_check_limit(ptype_ctx*, uint64_t max_iterations);

/* Order-independent 64-bit checksum of n reply records (sum over j of
 * mix64(record j, j) mod 2^64) — compares whole streams across GPUs and against
 * a CPU reference without moving the replies.  replies: device pointer when
 * PTYPE_F_DEVICE_PTRS, else host. */
int ptype_b200_reply_checksum(ptype_ctx*, int dev, const ptype_reply* replies, size_t n,
                              uint64_t index_base, uint64_t* checksum, uint32_t flags);

/* The same checksum for hosts that cannot allocate device memory while they run.
 * (No counterpart in the reference: the checksum itself is an addition of this
 * library, and so is this variant.)
 *
 * Who needs it: a host whose dispatch loop must never hit cudaMalloc/cudaFree,
 * because those calls hold up every stream of the GPU while they run, and a host
 * that has already given all device memory to its own pool or runs under a fixed
 * quota, where a late cudaMalloc can simply fail.  ptype_b200_reply_checksum
 * allocates its partial-sum buffer on first use; this entry point never calls
 * cudaMalloc, cudaFree or any allocating routine.  It needs no ptype_ctx.
 *
 * The partial sums go to a 128 KiB array compiled into the library, present on
 * every GPU that loads it until the process exits.  That fixes the maximum:
 * n <= 67 108 864 replies (16384 partial sums x 4096 replies each, 1 GiB of reply
 * records).  A larger n is refused with PTYPE_E_LIMIT and a message in
 * ptype_b200_last_error(); split the range (checksums of sub-ranges add up when
 * each is given its own index_base) or use ptype_b200_reply_checksum.
 *
 * device      CUDA ordinal the pointers belong to (made current for the call, then
 *             the caller's device is restored).
 * d_replies   device pointer, n records;  d_checksum  device pointer, one uint64.
 * cuda_stream a cudaStream_t (NULL = default stream); the two kernels are enqueued
 *             there and the call returns without waiting.
 *
 * The first call on each GPU also looks up where that GPU keeps the array; make
 * that first call before starting any stream capture.  All calls on a GPU share the
 * one array: two calls whose kernels can overlap (different streams, or several
 * threads) race on it and both results are garbage — keep calls on one stream or
 * order them yourself.  Nothing in the array has to be cleared between calls: each
 * call overwrites every entry it later reads. */
int ptype_b200_reply_checksum_noalloc(int device, const ptype_reply* d_replies, size_t n,
                                      uint64_t index_base, uint64_t* d_checksum, void* cuda_stream);

/* ---------------------------------------------- formats either side of the path */
/* All pointers below are DEVICE pointers on `dev`; the calls wait for completion.
 *
 * resolve_names: "Receiver.Method" strings -> hashes, split at the last '.' like net/rpc's
 * server (readRequestHeader [Go 1.13 net/rpc]).  names holds n NUL-padded strings of `stride`
 * bytes each (stride a multiple of 16; a string that fills its slot needs no NUL).  out[j]
 * (may be NULL) receives the pair; patch (may be NULL) gets reqs[j].method/.recv written. */
int ptype_b200_resolve_names(ptype_ctx*, int dev, const char* names, uint32_t stride, size_t n,
                             ptype_method_ref* out, ptype_req* patch);
/* gob_encode_replies: the reply VALUES as encoding/gob messages, `LL 04 00 <int>` each, in
 * call order, nothing for replies whose status is not OK (their body is net/rpc's empty
 * invalidRequest).  offsets[0..n] receive the byte position of every message (offsets[n] =
 * total); out receives the bytes that fit in out_cap; *total_bytes the size needed. */
int ptype_b200_gob_encode_replies(ptype_ctx*, int dev, const ptype_reply* replies, size_t n,
                                  uint64_t* offsets, uint8_t* out, size_t out_cap, uint64_t* total_bytes);
/* gob_decode_args: n gob struct bodies of calculator.Args / prime.Args (field delta, gob int)*
 * 00, zero fields omitted; body j spans bytes[offsets[j] .. offsets[j+1]).  args receives
 * args_per_call int64 per call, status[j] (may be NULL) 0 = ok, 1 = malformed. */
int ptype_b200_gob_decode_args(ptype_ctx*, int dev, const uint8_t* bytes, const uint64_t* offsets, size_t n,
                               int64_t* args, int args_per_call, uint8_t* status);

/* ---- net/rpc's byte streams on the device -------------------------------------
 * What travels on a connection dialled by rpc.DialHTTP (cluster/rpc.go:277) after the CONNECT
 * handshake: encoding/gob messages - per call an rpc.Request{ServiceMethod, Seq} value and an args
 * value from the client (cluster/rpc.go:65,88), an rpc.Response{ServiceMethod, Seq, Error} value
 * and a reply value from the server (rpc.Register/HandleHTTP, example/calculator/server/
 * server.go:16-20), with the type definitions gob sends before a type's first use.
 *
 * wire_decode takes the client-to-server bytes of n_streams connections (stream s spans
 * bytes[stream_off[s] .. stream_off[s+1]), each starting at a message boundary; it may begin with
 * the connection's type definitions or in mid-conversation) and writes one ptype_req per call:
 * method/recv hashed from the ServiceMethod string (split at the last '.'), a0..a2 from the args
 * value, svc_key and seq from the stream's descriptor (call i of stream s gets seq0 + i*seq_step:
 * the connection IS the client's routing decision, so the sequence number is the caller's to
 * state).  calls[] keeps what the answer must echo: net/rpc's own Seq and where the method string
 * lies in `bytes`.  call_off[0..n_streams] receives the index of every stream's first call.
 * *total_calls is always set; PTYPE_E_LIMIT (nothing decoded) when it exceeds cap.
 *
 * wire_encode writes the server-to-client bytes for those calls from their reply records:
 * Response header + `int` reply, or, for a status other than OK, the Error text net/rpc's server
 * sends ("rpc: can't find service X", ...) + the empty struct.  state[s] (one byte per stream,
 * 0 for a fresh connection, updated in place) remembers which definitions the connection has
 * seen: bit 0 rpc.Response, bit 1 the empty struct.  response_id / empty_id are the gob type ids
 * the sending process assigned to those two types (Go numbers user types from 65 in order of
 * first use).  out_off[0..n_streams] receives where every stream's bytes start in `out`;
 * *total_bytes is always set; PTYPE_E_LIMIT (nothing written, state untouched) beyond out_cap.
 * All pointers are device pointers on `dev` (`bytes` 16-byte aligned, as cudaMalloc gives it); both
 * calls wait for completion. */
#define PTYPE_WIRE_MALFORMED 1u
typedef struct ptype_wire_stream {
    uint64_t svc_key;  /* as ptype_req.svc_key: the service this server's node belongs to */
    uint64_t seq0;     /* ptype_req.seq of the stream's first call ...                      */
    uint64_t seq_step; /* ... and the step to the next one                                  */
} ptype_wire_stream;
typedef struct ptype_wire_call {
    uint64_t rpc_seq; /* rpc.Request.Seq: net/rpc's per-connection call number, echoed back */
    uint64_t sm_off;  /* the ServiceMethod string: offset into the request bytes ...         */
    uint32_t sm_len;  /* ... and length                                                       */
    uint32_t flags;   /* PTYPE_WIRE_MALFORMED: the call's messages did not parse             */
} ptype_wire_call;
int ptype_b200_wire_decode(ptype_ctx*, int dev, const uint8_t* bytes, const uint64_t* stream_off, size_t n_streams,
                           const ptype_wire_stream* streams, uint64_t* call_off, ptype_req* reqs,
                           ptype_wire_call* calls, size_t cap, uint64_t* total_calls);
int ptype_b200_wire_encode(ptype_ctx*, int dev, const uint8_t* request_bytes, const uint64_t* call_off,
                           size_t n_streams, const ptype_wire_call* calls, const ptype_reply* replies,
                           uint32_t response_id, uint32_t empty_id, uint8_t* state, uint64_t* out_off, uint8_t* out,
                           size_t out_cap, uint64_t* total_bytes);

/* ---- the balancer's counters on the device ---------------------------------
 * Every reference Client owns one counter, `seq` (cluster/rpc.go:129); roundRobinSelect takes its
 * next value for each call that finds a non-empty client list - the length check at rpc.go:178-180
 * comes before atomic.AddUint64 at rpc.go:181, so calls to a service without nodes, like calls to
 * a service no Client exists for, consume nothing.  A context keeps those counters on the device,
 * one per service of the installed image, all starting at 0 like a fresh Client's.
 *
 * assign_seq fills ptype_req.seq of n device-resident records exactly as n sequential
 * Client.Call()s in batch order would have drawn them: record j gets
 *   counter[service_j] + 1 + (number of earlier counted records of the same service),
 * the counters move past the batch, and records that consume nothing get seq 0.  The host then
 * need not carry seq at all (nor keep per-service counters for a batch that mixes services).
 * Counters follow their services, matched by key, through publish_image / broadcast_image /
 * update_service; a service missing from a new image loses its counter, and an image built under
 * a different key salt (only after a 64-bit key collision among the registered names) starts all
 * of them at 0 - seq_get/seq_set carry them over by name in that case.
 * Device pointers only; PTYPE_F_ASYNC is honoured.  PTYPE_E_LIMIT: tables beyond 51 200 slots
 * (the per-CTA histogram is one 32-bit count per slot in shared memory). */
int ptype_b200_assign_seq(ptype_ctx*, int dev, ptype_req* d_reqs, size_t n, uint32_t flags);
/* The same draw, with the numbers written to a plain array d_seq[0..n) instead of into the records
 * (which are only read), and the dispatch that takes them from there: call j is routed with
 * d_seq[j], its record's seq field is ignored.  Same replies as assign_seq + dispatch, for less
 * memory traffic: storing 8 bytes into every 48-byte record costs the memory system a read and a
 * write of a 32-byte sector per record, an array of 8-byte numbers is written and read once.
 * Device pointers only; PTYPE_F_ASYNC is honoured; dispatch_seq accepts PTYPE_F_GLOBAL_TABLE. */
int ptype_b200_assign_seq_array(ptype_ctx*, int dev, const ptype_req* d_reqs, size_t n, uint64_t* d_seq,
                                uint32_t flags);
int ptype_b200_dispatch_seq(ptype_ctx*, int dev, const ptype_req* d_reqs, const uint64_t* d_seq, size_t n,
                            ptype_reply* d_replies, uint32_t flags);
/* Read / set one service's counter (the value the next call will exceed by one).
 * PTYPE_E_NOT_FOUND: the installed image has no such service. */
int ptype_b200_seq_get(ptype_ctx*, int dev, const char* service, uint64_t* value);
int ptype_b200_seq_set(ptype_ctx*, int dev, const char* service, uint64_t value);

/* Kernel launches issued by this context so far (bench bookkeeping). */
int ptype_b200_launch_count(ptype_ctx*, uint64_t* launches);

#ifdef __cplusplus
}
#endif
#endif /* PTYPE_B200_H */
