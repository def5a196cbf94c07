/*
 * ptype_oracle.h — CPU restatement of the reference's batched-dispatch path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product: only
 * tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may build, load or call it, and only as the checker or the timed CPU
 * baseline.  The product path (ptype_b200/) never links or imports it.
 *
 * Parity pinning: the reference is Go and no Go toolchain exists in this image,
 * so oracle/_ref cannot be built.  The restatement is pinned instead against
 * every result the reference's own tests fix for this path (SURVEY.md §8c,
 * G1..G10; see tests/test_oracle_golden.py and oracle/selftest.c).  Multiply,
 * Add and Check have no reference tests: they are pinned only by the Go
 * language rules (wrapping int64, truncated %), i.e. PARITY UNPINNED for those
 * three bodies beyond the 7*8 example in example/calculator/client/client.go:38-44.
 *
 * Every function cites the reference lines it follows.  Record layouts come
 * from include/ptype_b200.h (the interface), nothing else is shared.
 */
#ifndef PTYPE_ORACLE_H
#define PTYPE_ORACLE_H

#include <stddef.h>
#include <stdint.h>

#include "../include/ptype_b200.h"

#ifdef __cplusplus
extern "C" {
#endif

/* ---- Go stdlib pieces the path calls (hash/fnv, strconv, path/filepath) ---- */
uint32_t po_fnv1a32(const void* p, size_t n);                 /* fnv.New32a, rpc.go:267 */
uint64_t po_fnv1a64_salted(const void* p, size_t n, uint64_t salt);
int po_itoa(int64_t v, char* out);                            /* strconv.Itoa, rpc.go:268 */
/* filepath.Join(elems...) + "/"  (registry.go:168-170); returns length or -1 */
int po_etcd_key(char* out, size_t cap, const char* const* elems, int n);

/* ---- an ordered KV standing in for etcd (sorted range, put, delete) ---- */
typedef struct po_kv po_kv;
po_kv* po_kv_new(void);
void po_kv_free(po_kv*);
void po_kv_put(po_kv*, const char* key, const char* value);
int po_kv_delete(po_kv*, const char* key); /* number deleted */
/* clientv3 Get(prefix, WithPrefix, SortByKey Ascend) (registry.go:88-91): returns
 * count, and the i-th key/value through po_kv_at on [first, first+count). */
size_t po_kv_range_prefix(const po_kv*, const char* prefix, size_t* first);
const char* po_kv_key_at(const po_kv*, size_t i);
const char* po_kv_value_at(const po_kv*, size_t i);
size_t po_kv_len(const po_kv*);

/* ---- cluster.Node and its JSON (registry.go:23-26,52-53,102,161) ---- */
typedef struct po_node {
    char address[256];
    int64_t port;
} po_node;
int po_node_marshal(const po_node*, char* out, size_t cap);
/* json.Unmarshal into an existing Node: fields absent from the JSON keep their
 * previous value (this is what makes registry.go:99,102 reuse observable). 0 = ok. */
int po_node_unmarshal(const char* json, po_node* inout);

/* ---- registry (registry.go:51-86, 93-117, 152-166) ---- */
int po_register(po_kv*, const char* service, const char* node_name, const char* host, int64_t port);
typedef struct po_service_list {
    char name[256];
    po_node* nodes;
    size_t n;
} po_service_list;
/* Services(): map service -> nodes in key order.  Output sorted by first
 * appearance in the range (Go map order is unobservable).  Returns count or -1. */
long po_services(const po_kv*, po_service_list** out);
void po_services_free(po_service_list*, size_t n);
/* nodes(serviceName): range services/<svc>/ ; returns count or -1 */
long po_nodes(const po_kv*, const char* service, po_node** out);

/* ---- balancer (rpc.go:172-183, 252-270) ---- */
/* hashConnectionIndex: int(FNV-1a-32(localAddr+Itoa(connNumber))) % nodeSize */
int64_t po_hash_connection_index(const char* local_addr, int64_t conn_number, int64_t node_size);
/* selectNodes: writes indices into nodes[] (0-based) of the selected list; returns k.
 * max_connections < 0 is rejected (-1): Go would panic in make() (rpc.go:257). */
long po_select_nodes(const char* local_addr, size_t n_nodes, int64_t max_connections,
                     uint32_t* selected_idx, size_t cap);
/* roundRobinSelect: seq = value AFTER the atomic add. Returns index >= 0,
 * -1 when the list is empty (nil client), -2 when Go would panic (negative index). */
int64_t po_round_robin(uint64_t seq_after_add, size_t n_clients);

/* ---- method bodies ---- */
int64_t po_multiply(int64_t a, int64_t b); /* calculator.go:9-12 */
int64_t po_add(int64_t a, int64_t b);      /* defined by the build */
int64_t po_check(int64_t min, int64_t max, int64_t target); /* prime.go:15-32, no sleep */
uint64_t po_check_width(int64_t min, int64_t max, int64_t target); /* loop trip bound */

/* ---- optimus stream (coordinator.go:67-73) ---- */
/* splitWork(target): emits (min,max,target) triples; returns count. out may be NULL. */
size_t po_split_work(int64_t target, int64_t (*out)[3], size_t cap);

/* ---- a whole-registry snapshot and batched dispatch over packed records ---- */
typedef struct po_snapshot po_snapshot;
/* Built from the KV the way a host would: Services() + one Client per service
 * with (local_addr, max_connections).  `declared` lists services that have a
 * Client but no nodes (may be NULL). */
po_snapshot* po_snapshot_build(const po_kv*, const char* local_addr, int64_t max_connections,
                               const char* const* declared, size_t n_declared, uint64_t key_salt);
void po_snapshot_free(po_snapshot*);
/* rpc.Register on the nodes of `service`: receiver is "Calculator" or "Prime". */
int po_snapshot_expose(po_snapshot*, const char* service, const char* receiver);
size_t po_snapshot_n_nodes(const po_snapshot*);
/* getSelectedNodes() of the Client for `service` as global node ids. */
long po_snapshot_selected(const po_snapshot*, const char* service, uint32_t* out, size_t cap);
int po_snapshot_node(const po_snapshot*, uint32_t node_id, po_node* out, char* service, size_t cap);

/* One Client.Call on strings (rpc.go:59-67 + net/rpc server resolution). */
void po_call(const po_snapshot*, const char* service, uint64_t seq_after_add,
             const char* service_method, const int64_t args[3], uint64_t check_limit,
             ptype_reply* reply);
/* The same over packed records. */
void po_dispatch(const po_snapshot*, const ptype_req* reqs, size_t n, ptype_reply* replies,
                 uint64_t check_limit);
void po_dispatch_mt(const po_snapshot*, const ptype_req* reqs, size_t n, ptype_reply* replies,
                    uint64_t check_limit, int threads);
/* The counters behind the seq field: n sequential Client.Call()s in batch order, each on the
 * Client of its record's service (one Client, one counter: rpc.go:129).  A call that finds its
 * Client's list non-empty takes the counter's next value (rpc.go:177-182: the length check comes
 * first, so an empty list - like a service no Client exists for - consumes nothing; such records
 * get seq 0).  counters[i] belongs to the i-th Client in key order (po_snapshot_client_index) and
 * is advanced. */
void po_assign_seq(const po_snapshot*, ptype_req* reqs, size_t n, uint64_t* counters);
size_t po_snapshot_n_clients(const po_snapshot*);
long po_snapshot_client_index(const po_snapshot*, const char* service); /* -1: no such Client */
uint64_t po_reply_checksum(const ptype_reply* replies, size_t n, uint64_t index_base);
void po_method_key(const char* service_method, uint32_t* method, uint32_t* recv);

/* ---- encoding/gob value encodings ---- */
int po_gob_uint(uint64_t v, uint8_t* out);
int po_gob_int(int64_t v, uint8_t* out);

/* ---- config 1: the reference's real path over loopback TCP (oracle/loopback.c) ---- */
typedef struct po_loopback_stats {
    double seconds;
    uint64_t calls;
    uint64_t bytes_client_to_server;
    uint64_t bytes_server_to_client;
} po_loopback_stats;
/* n sequential synchronous Client.Call()s of one Client holding k connections; every node
 * registered `receivers` (bit0 Calculator, bit1 Prime).  replies[j].node = connection index. */
int po_loopback_run(const ptype_req* reqs, size_t n, ptype_reply* replies, int k, uint32_t receivers,
                    uint64_t check_limit, po_loopback_stats* stats);
int po_wire_streams(const ptype_req* reqs, size_t n, uint32_t receivers, uint64_t check_limit, uint8_t* c2s,
                    size_t c2s_cap, uint64_t* c2s_off, uint8_t* s2c, size_t s2c_cap, uint64_t* s2c_off);
int po_gob_reply_int_message(int64_t v, uint8_t* out);
int po_gob_args_body(const int64_t* a, int nargs, uint8_t* out);
int po_gob_request_typedef(uint8_t* out);

#ifdef __cplusplus
}
#endif
#endif
