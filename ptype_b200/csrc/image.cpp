// image.cpp — host-side registry image builder and readers (no GPU needed).
//
// The image is the device form of what Registry.Services() returns
// (cluster/registry.go:93-117) together with the two inputs of the selection rule,
// localAddr and ConnConfig.MaxConnections (cluster/rpc.go:19-38,128,252-270).
// Node order inside a service and the global node numbering follow the etcd
// range order the reference reads: ascending byte order of the keys
// services/<service>/<node>/ (cluster/registry.go:64,88-91,168-170).
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <new>
#include <string>
#include <utility>
#include <vector>

#include "../../include/ptype_b200.h"
#include "table_hash.h"

static_assert(sizeof(ptype_image_header) == 192, "image header layout");
static_assert(sizeof(ptype_req) == 48 && sizeof(ptype_reply) == 16 && sizeof(ptype_slot) == 16, "record layout");

namespace {

struct NodeRec {
    std::string address;
    int port = 0;
};

struct ServiceRec {
    std::map<std::string, NodeRec> nodes;  // by node name
    uint32_t recv_mask = 0;
    uint32_t order_index = 0;  // position in Placement::order while the placement is valid
};

struct KnownMethod {
    const char* recv;
    uint32_t recv_bit;
    const char* method;
    uint32_t opcode;
};

// Exported methods of the example receivers: Calculator.Multiply
// (example/calculator/calculator.go:9), Prime.Check (example/optimus/prime.go:15),
// and Calculator.Add, which the build defines.
const KnownMethod kKnownMethods[] = {
    {"Calculator", 0, "Multiply", PTYPE_OP_MULTIPLY},
    {"Calculator", 0, "Add", PTYPE_OP_ADD},
    {"Prime", 1, "Check", PTYPE_OP_CHECK},
};

// Largest slot array the builder sizes for staging in shared memory: 192 KiB.  With the method
// table and hashes the CTA then stays inside 196 KiB, the largest carve-out that leaves the SM
// 60 KiB of L1 for the request loads in flight (a 200 KiB table leaves 28 KiB and cost C3 10 %,
// see plan_dispatch in dispatch.cu).
constexpr uint32_t kSmemSlotBudget = 192 * 1024 / sizeof(ptype_slot);

bool valid_name(const char* s) {
    if (!s || !*s) return false;
    if (!std::strcmp(s, ".") || !std::strcmp(s, "..")) return false;  // filepath.Join would fold these
    return std::strchr(s, '/') == nullptr;  // would change Split(key,"/")[1] (registry.go:106-108)
}

size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

}  // namespace

// What finish() derives from the SET of service names alone: key order, salt, table size and
// the cuckoo placement.  A node joining or leaving an existing service (the common live update,
// registry.go:119-150) changes none of it, so it is kept between finish() calls and dropped
// whenever a service name is added or the table size is forced.
struct Placement {
    bool valid = false;
    std::vector<const std::pair<const std::string, ServiceRec>*> order;  // etcd key order
    std::vector<uint64_t> keys;
    uint64_t salt = 0;
    uint32_t n_slots = 0;
    uint32_t table_seed = 0;
    std::vector<uint32_t> slot_index;  // per service (in `order`), its slot in the table
    std::vector<uint32_t> by_key;      // `order` indices sorted by key (host-side binary search)
    // Dense copies of each service's node count and receiver mask, kept current by the
    // mutators, so finish() walks two flat arrays instead of 10 000 scattered map nodes.
    std::vector<uint32_t> n_nodes, recv_mask;
    uint64_t total_nodes = 0;

    void note(const ServiceRec& rec) {
        if (!valid) return;
        total_nodes += rec.nodes.size();
        total_nodes -= n_nodes[rec.order_index];
        n_nodes[rec.order_index] = static_cast<uint32_t>(rec.nodes.size());
        recv_mask[rec.order_index] = rec.recv_mask;
    }
};

struct ptype_builder {
    std::string local_addr;
    int max_connections = 0;
    uint32_t forced_slots = 0;
    bool device_only = false;
    std::map<std::string, ServiceRec> services;  // node-stable: Placement points into it
    Placement placement;
};

// Key order, salt, table size and cuckoo placement for the builder's current set of service names.
static int compute_placement(ptype_builder* b) {
    Placement& pl = b->placement;
    pl.valid = false;
    const size_t n_services = b->services.size();

    // Key order of the etcd range: compare the full keys bytewise.  '/' (0x2f) sorts
    // after '-' and '.', so "foo-x" precedes "foo" here although the names alone
    // would sort the other way.
    std::vector<std::pair<std::string, const std::pair<const std::string, ServiceRec>*>> order;
    order.reserve(n_services);
    for (const auto& kv : b->services) order.emplace_back("services/" + kv.first + "/", &kv);
    std::sort(order.begin(), order.end(),
              [](const auto& a, const auto& c) { return a.first < c.first; });
    pl.order.resize(n_services);
    pl.n_nodes.resize(n_services);
    pl.recv_mask.resize(n_services);
    pl.total_nodes = 0;
    for (size_t i = 0; i < n_services; ++i) {
        pl.order[i] = order[i].second;
        ServiceRec& rec = b->services.find(order[i].second->first)->second;
        rec.order_index = static_cast<uint32_t>(i);
        pl.n_nodes[i] = static_cast<uint32_t>(rec.nodes.size());
        pl.recv_mask[i] = rec.recv_mask;
        pl.total_nodes += rec.nodes.size();
    }

    // Table size: half full when that fits the shared-memory budget, otherwise
    // fill the budget (up to 85 % load: 3-way cuckoo places that reliably), otherwise
    // a half-full table probed in L2.
    if (b->forced_slots) {
        // 3-way cuckoo hashing stops placing reliably a little above 90 % load
        if (static_cast<uint64_t>(b->forced_slots) * 9 < static_cast<uint64_t>(n_services) * 10 ||
            b->forced_slots <= n_services)
            return PTYPE_E_INVAL;
        pl.n_slots = b->forced_slots;
    } else {
        size_t want = std::max<size_t>(8, align_up(2 * n_services, 8));
        if (want > kSmemSlotBudget && n_services * 20 <= static_cast<size_t>(kSmemSlotBudget) * 17)
            want = kSmemSlotBudget;
        pl.n_slots = static_cast<uint32_t>(want);
    }

    // Pick a salt under which all registered names have distinct 64-bit keys (H7) AND the cuckoo
    // placement settles.  The three candidate slots of a key depend on its 32-bit fold only, so
    // four names whose keys fold alike can never be placed under any table seed; a new salt gives
    // them new keys (and folds), which is why a placement failure moves on to the next salt.
    constexpr uint32_t kFree = 0xFFFFFFFFu;
    constexpr uint32_t kSeedsPerSalt = 8;
    std::vector<uint32_t> table;
    pl.keys.assign(n_services, 0);
    bool placed_all = false;
    for (pl.salt = 0; !placed_all; ++pl.salt) {
        if (pl.salt == 64) return PTYPE_E_COLLISION;
        for (size_t i = 0; i < n_services; ++i)
            pl.keys[i] = ptype::fnv1a64(pl.order[i]->first.data(), pl.order[i]->first.size(), pl.salt);
        std::vector<uint64_t> sorted(pl.keys);
        std::sort(sorted.begin(), sorted.end());
        if (std::adjacent_find(sorted.begin(), sorted.end()) != sorted.end()) continue;

        // Random-walk insertion; a walk that does not settle restarts the whole table under the
        // next seed.  `table` holds service indices while placing.
        for (pl.table_seed = 0; pl.table_seed < kSeedsPerSalt && !placed_all; ++pl.table_seed) {
            table.assign(pl.n_slots, kFree);
            uint64_t rng = 0x9E3779B97F4A7C15ull ^ pl.table_seed;
            bool ok = true;
            for (size_t i = 0; i < n_services && ok; ++i) {
                uint32_t cur = static_cast<uint32_t>(i);
                ok = false;
                for (int kick = 0; kick < 2000; ++kick) {
                    const uint64_t key = pl.keys[cur];
                    const uint32_t lo = static_cast<uint32_t>(key), hi = static_cast<uint32_t>(key >> 32);
                    uint32_t cand[PTYPE_SLOT_WAYS];
                    bool placed = false;
                    for (int w = 0; w < PTYPE_SLOT_WAYS; ++w) {
                        cand[w] = ptype::slot_of(lo, hi, pl.table_seed, w, pl.n_slots);
                        if (table[cand[w]] == kFree) {
                            table[cand[w]] = cur;
                            placed = true;
                            break;
                        }
                    }
                    if (placed) {
                        ok = true;
                        break;
                    }
                    rng = rng * 6364136223846793005ull + 1442695040888963407ull;
                    const uint32_t victim = cand[(rng >> 33) % PTYPE_SLOT_WAYS];
                    std::swap(cur, table[victim]);
                }
            }
            placed_all = ok;
        }
        if (placed_all) break;  // keep pl.salt; pl.table_seed is one past the seed that worked
    }
    pl.table_seed -= 1;
    pl.slot_index.assign(n_services, 0);
    for (uint32_t s = 0; s < pl.n_slots; ++s)
        if (table[s] != kFree) pl.slot_index[table[s]] = s;

    pl.by_key.resize(n_services);
    for (size_t i = 0; i < n_services; ++i) pl.by_key[i] = static_cast<uint32_t>(i);
    std::sort(pl.by_key.begin(), pl.by_key.end(),
              [&](uint32_t a, uint32_t c) { return pl.keys[a] < pl.keys[c]; });
    pl.valid = true;
    return PTYPE_OK;
}

extern "C" {

uint32_t ptype_b200_fnv1a32(const void* data, size_t len) { return ptype::fnv1a32(data, len); }
uint64_t ptype_b200_fnv1a64(const void* data, size_t len) { return ptype::fnv1a64(data, len, 0); }

uint64_t ptype_b200_service_key(const char* service_name, uint64_t salt) {
    return ptype::fnv1a64(service_name, std::strlen(service_name), salt);
}

void ptype_b200_method_key(const char* service_method, uint32_t* method, uint32_t* recv) {
    size_t len = std::strlen(service_method);
    *method = ptype::fnv1a32(service_method, len);
    const char* dot = std::strrchr(service_method, '.');
    *recv = dot ? ptype::fnv1a32(service_method, static_cast<size_t>(dot - service_method))
                : PTYPE_RECV_ILL_FORMED;
}

// FNV-1a-32(localAddr + strconv.Itoa(connNumber)) (cluster/rpc.go:267-269).
uint32_t ptype_b200_conn_hash(const char* local_addr, int conn_number) {
    std::string s(local_addr ? local_addr : "");
    s += std::to_string(conn_number);
    return ptype::fnv1a32(s.data(), s.size());
}

int ptype_b200_builder_new(const char* local_addr, int max_connections, ptype_builder** out) {
    if (!out || max_connections < 0) return PTYPE_E_INVAL;  // Go: make([]Node,0,neg) panics (rpc.go:257)
    if (local_addr && std::strlen(local_addr) >= PTYPE_LOCAL_ADDR_MAX) return PTYPE_E_LIMIT;
    ptype_builder* b = new (std::nothrow) ptype_builder();
    if (!b) return PTYPE_E_NOMEM;
    b->local_addr = local_addr ? local_addr : "";
    b->max_connections = max_connections;
    *out = b;
    return PTYPE_OK;
}

int ptype_b200_builder_set_slots(ptype_builder* b, uint32_t n_slots) {
    if (!b) return PTYPE_E_INVAL;
    b->forced_slots = n_slots;
    b->placement.valid = false;
    return PTYPE_OK;
}

int ptype_b200_builder_register(ptype_builder* b, const char* service, const char* node_name,
                                const char* address, int port) {
    if (!b || !valid_name(service) || !valid_name(node_name) || !address) return PTYPE_E_INVAL;
    auto it = b->services.find(service);
    if (it == b->services.end()) {
        it = b->services.emplace(service, ServiceRec()).first;
        b->placement.valid = false;
    }
    NodeRec& n = it->second.nodes[node_name];  // Put on an existing key overwrites
    n.address = address;
    n.port = port;
    b->placement.note(it->second);
    return PTYPE_OK;
}

// Deleting the key services/<service>/<node_name>/ (lease expiry, registry.go:58-83, or an
// explicit delete as in registry_test.go:179-191).  The service stays known with the nodes left.
int ptype_b200_builder_unregister(ptype_builder* b, const char* service, const char* node_name) {
    if (!b || !valid_name(service) || !valid_name(node_name)) return PTYPE_E_INVAL;
    auto it = b->services.find(service);
    if (it == b->services.end() || it->second.nodes.erase(node_name) == 0) return PTYPE_E_NOT_FOUND;
    b->placement.note(it->second);
    return PTYPE_OK;
}

int ptype_b200_builder_set_device_only(ptype_builder* b, int on) {
    if (!b) return PTYPE_E_INVAL;
    b->device_only = on != 0;
    return PTYPE_OK;
}

int ptype_b200_builder_declare(ptype_builder* b, const char* service) {
    if (!b || !valid_name(service)) return PTYPE_E_INVAL;
    if (b->services.find(service) == b->services.end()) {
        b->services[service];
        b->placement.valid = false;
    }
    return PTYPE_OK;
}

int ptype_b200_builder_expose(ptype_builder* b, const char* service, const char* receiver) {
    if (!b || !valid_name(service) || !receiver) return PTYPE_E_INVAL;
    for (const KnownMethod& m : kKnownMethods) {
        if (!std::strcmp(m.recv, receiver)) {
            if (b->services.find(service) == b->services.end()) b->placement.valid = false;
            ServiceRec& rec = b->services[service];
            rec.recv_mask |= 1u << m.recv_bit;
            b->placement.note(rec);
            return PTYPE_OK;
        }
    }
    return PTYPE_E_NOT_FOUND;
}

void ptype_b200_builder_free(ptype_builder* b) { delete b; }
void ptype_b200_image_free(void* image) { std::free(image); }

int ptype_b200_builder_finish(ptype_builder* b, uint64_t epoch, void** image, size_t* bytes) {
    if (!b || !image || !bytes) return PTYPE_E_INVAL;
    const size_t n_services = b->services.size();
    if (n_services >= 0x7FFFFFFFu) return PTYPE_E_LIMIT;
    if (!b->placement.valid) {
        const int rc = compute_placement(b);
        if (rc != PTYPE_OK) return rc;
    }
    const Placement& pl = b->placement;
    const size_t n_nodes = pl.total_nodes;
    if (n_nodes >= PTYPE_NO_NODE) return PTYPE_E_LIMIT;
    for (size_t i = 0; i < n_services; ++i)
        if (pl.n_nodes[i] > 0xFFFFFFu) return PTYPE_E_LIMIT;
    const uint64_t salt = pl.salt;
    const uint32_t n_slots = pl.n_slots, table_seed = pl.table_seed;

    const uint32_t n_conn = static_cast<uint32_t>(b->max_connections);
    const uint32_t n_methods = sizeof(kKnownMethods) / sizeof(kKnownMethods[0]);

    // String pool.
    std::string pool;
    pool.push_back('\0');  // offset 0 = ""
    auto intern = [&pool](const std::string& s) {
        uint32_t off = static_cast<uint32_t>(pool.size());
        pool.append(s);
        pool.push_back('\0');
        return off;
    };

    size_t off = sizeof(ptype_image_header);
    const size_t off_slots = off;
    off += static_cast<size_t>(n_slots) * sizeof(ptype_slot);
    const size_t off_conn = off;
    off = align_up(off + static_cast<size_t>(n_conn) * sizeof(uint32_t), 16);
    const size_t off_methods = off;
    off += n_methods * sizeof(ptype_method);
    const size_t off_mtable = off;
    off += PTYPE_MTABLE_SLOTS * sizeof(ptype_method);
    const size_t off_order = off;
    off = align_up(off + static_cast<size_t>(n_slots) * sizeof(uint32_t), 16);
    const size_t device_bytes = off;
    const bool host_sections = !b->device_only;
    const size_t off_services = off;
    if (host_sections) off += n_services * sizeof(ptype_service);
    const size_t off_nodes = off;
    if (host_sections) off += n_nodes * sizeof(ptype_node);
    const size_t off_strings = off;

    std::vector<ptype_service> services(n_services);
    std::vector<ptype_node> nodes(host_sections ? n_nodes : 0);
    uint32_t next_node = 0;
    for (size_t i = 0; i < n_services; ++i) {
        ptype_service& s = services[i];
        s.key = pl.keys[i];
        s.name_off = 0;
        s.node_base = next_node;
        s.n_nodes = pl.n_nodes[i];
        s.recv_mask = pl.recv_mask[i];
        if (!host_sections) {
            next_node += s.n_nodes;
            continue;
        }
        const std::string& name = pl.order[i]->first;
        const ServiceRec& rec = pl.order[i]->second;
        s.name_off = intern(name);
        // Order inside the service = byte order of the keys services/<svc>/<node>/, i.e. of the
        // node names WITH their trailing '/': "node10" before "node2", and "a-b" before "a"
        // because '-' sorts below '/'.
        std::vector<std::pair<std::string, const std::pair<const std::string, NodeRec>*>> by_key;
        by_key.reserve(rec.nodes.size());
        for (const auto& nk : rec.nodes) by_key.emplace_back(nk.first + "/", &nk);
        std::sort(by_key.begin(), by_key.end(), [](const auto& a, const auto& c) { return a.first < c.first; });
        for (const auto& e : by_key) {
            const auto& nk = *e.second;
            ptype_node& nd = nodes[next_node++];
            nd.address_off = intern(nk.second.address);
            nd.name_off = intern(nk.first);
            nd.service = static_cast<uint32_t>(i);
            nd.port = nk.second.port;
        }
    }
    const size_t total = align_up(off_strings + pool.size(), 16);
    if (total > 0xFFFFFFFFull) return PTYPE_E_LIMIT;

    unsigned char* blob = static_cast<unsigned char*>(std::calloc(1, total));
    if (!blob) return PTYPE_E_NOMEM;

    // Every key sits in one of its three candidate slots (compute_placement), so a device
    // lookup is three independent probes at most.
    ptype_slot* slots = reinterpret_cast<ptype_slot*>(blob + off_slots);
    for (uint32_t i = 0; i < n_slots; ++i) slots[i].node_base = PTYPE_SLOT_EMPTY;
    for (size_t i = 0; i < n_services; ++i) {
        ptype_slot& e = slots[pl.slot_index[i]];
        e.key = services[i].key;
        e.node_base = services[i].node_base;
        e.n_and_recv = services[i].n_nodes | (services[i].recv_mask << 24);
    }
    // Key-order rank per slot: global node ids are ranks in the sorted key range, so a node joining
    // or leaving service r moves the ids of exactly the services whose rank exceeds r - including
    // node-less ones, which node_base alone cannot order (ptype_b200_update_service).
    uint32_t* order = reinterpret_cast<uint32_t*>(blob + off_order);
    for (uint32_t i = 0; i < n_slots; ++i) order[i] = PTYPE_SLOT_EMPTY;
    for (size_t i = 0; i < n_services; ++i) order[pl.slot_index[i]] = static_cast<uint32_t>(i);
    uint32_t* conn = reinterpret_cast<uint32_t*>(blob + off_conn);
    for (uint32_t i = 0; i < n_conn; ++i)
        conn[i] = ptype_b200_conn_hash(b->local_addr.c_str(), static_cast<int>(i));
    ptype_method* methods = reinterpret_cast<ptype_method*>(blob + off_methods);
    for (uint32_t i = 0; i < n_methods; ++i) {
        const KnownMethod& m = kKnownMethods[i];
        std::string full = std::string(m.recv) + "." + m.method;
        methods[i].method_hash = ptype::fnv1a32(full.data(), full.size());
        methods[i].recv_hash = ptype::fnv1a32(m.recv, std::strlen(m.recv));
        methods[i].opcode = m.opcode;
        methods[i].recv_bit = m.recv_bit;
    }
    // Perfect-hash method table: pick an odd multiplier under which the registered method
    // hashes fall into distinct entries, so the device resolves a method with one load.
    ptype_method* mtable = reinterpret_cast<ptype_method*>(blob + off_mtable);
    uint32_t method_mul = 1;
    for (;; method_mul += 2) {
        if (method_mul > (1u << 20)) {
            std::free(blob);
            return PTYPE_E_COLLISION;
        }
        std::memset(mtable, 0, PTYPE_MTABLE_SLOTS * sizeof(ptype_method));
        bool ok = true;
        for (uint32_t i = 0; i < n_methods && ok; ++i) {
            ptype_method& e = mtable[ptype::mtable_index(methods[i].method_hash, method_mul)];
            ok = e.opcode == PTYPE_OP_NONE;
            e = methods[i];
        }
        if (ok) break;
    }
    // Host-side service records sorted by key for binary search.
    if (host_sections) {
        const std::vector<uint32_t>& perm = pl.by_key;
        std::vector<uint32_t> inv(n_services);
        for (size_t i = 0; i < n_services; ++i) inv[perm[i]] = static_cast<uint32_t>(i);
        ptype_service* out_services = reinterpret_cast<ptype_service*>(blob + off_services);
        for (size_t i = 0; i < n_services; ++i) out_services[i] = services[perm[i]];
        ptype_node* out_nodes = reinterpret_cast<ptype_node*>(blob + off_nodes);
        for (size_t i = 0; i < n_nodes; ++i) {
            out_nodes[i] = nodes[i];
            out_nodes[i].service = inv[nodes[i].service];
        }
    }
    std::memcpy(blob + off_strings, pool.data(), pool.size());

    ptype_image_header* h = reinterpret_cast<ptype_image_header*>(blob);
    h->magic = PTYPE_IMAGE_MAGIC;
    h->version = PTYPE_IMAGE_VERSION;
    h->epoch = epoch;
    h->total_bytes = total;
    h->key_salt = salt;
    h->n_services = static_cast<uint32_t>(n_services);
    h->n_slots = n_slots;
    h->n_nodes = static_cast<uint32_t>(n_nodes);
    h->max_connections = n_conn;
    h->n_conn_hash = n_conn;
    h->n_methods = n_methods;
    h->n_receivers = 2;
    h->host_sections = host_sections ? 1u : 0u;
    h->off_slots = static_cast<uint32_t>(off_slots);
    h->off_conn_hash = static_cast<uint32_t>(off_conn);
    h->off_methods = static_cast<uint32_t>(off_methods);
    h->off_nodes = static_cast<uint32_t>(off_nodes);
    h->off_services = static_cast<uint32_t>(off_services);
    h->off_strings = static_cast<uint32_t>(off_strings);
    h->strings_bytes = static_cast<uint32_t>(pool.size());
    h->device_bytes = static_cast<uint32_t>(device_bytes);
    h->table_seed = table_seed;
    h->method_mul = method_mul;
    h->off_mtable = static_cast<uint32_t>(off_mtable);
    h->off_order = static_cast<uint32_t>(off_order);
    std::snprintf(h->local_addr, sizeof h->local_addr, "%s", b->local_addr.c_str());
    h->body_checksum = ptype::image_body_hash(blob + sizeof(ptype_image_header),
                                              total - sizeof(ptype_image_header));
    *image = blob;
    *bytes = total;
    return PTYPE_OK;
}

int ptype_b200_image_validate(const void* image, size_t bytes) {
    if (!image || bytes < sizeof(ptype_image_header)) return PTYPE_E_BAD_IMAGE;
    const ptype_image_header* h = static_cast<const ptype_image_header*>(image);
    if (h->magic != PTYPE_IMAGE_MAGIC || h->version != PTYPE_IMAGE_VERSION) return PTYPE_E_BAD_IMAGE;
    if (h->total_bytes != bytes || (bytes & 7u) || h->n_slots == 0 || h->n_services >= h->n_slots)
        return PTYPE_E_BAD_IMAGE;
    if (h->n_methods > PTYPE_MAX_METHODS || h->n_conn_hash != h->max_connections)
        return PTYPE_E_BAD_IMAGE;
    auto inside = [&](uint64_t off, uint64_t len) { return off >= sizeof(*h) && off + len <= bytes; };
    if ((h->off_slots & 15u) || (h->off_methods & 15u)) return PTYPE_E_BAD_IMAGE;
    if (!inside(h->off_slots, uint64_t(h->n_slots) * sizeof(ptype_slot)) ||
        !inside(h->off_conn_hash, uint64_t(h->n_conn_hash) * 4) ||
        !inside(h->off_methods, uint64_t(h->n_methods) * sizeof(ptype_method)) ||
        !inside(h->off_mtable, uint64_t(PTYPE_MTABLE_SLOTS) * sizeof(ptype_method)) ||
        (h->off_mtable & 15u) || !(h->method_mul & 1u) ||
        !inside(h->off_order, uint64_t(h->n_slots) * 4) || (h->off_order & 3u) ||
        h->device_bytes < h->off_order + uint64_t(h->n_slots) * 4 ||
        (h->host_sections && (!inside(h->off_services, uint64_t(h->n_services) * sizeof(ptype_service)) ||
                              !inside(h->off_nodes, uint64_t(h->n_nodes) * sizeof(ptype_node)))) ||
        !inside(h->off_strings, h->strings_bytes) || h->device_bytes > bytes ||
        h->device_bytes < h->off_mtable + PTYPE_MTABLE_SLOTS * sizeof(ptype_method))
        return PTYPE_E_BAD_IMAGE;
    if (h->local_addr[PTYPE_LOCAL_ADDR_MAX - 1] != 0) return PTYPE_E_BAD_IMAGE;
    const unsigned char* p = static_cast<const unsigned char*>(image);
    if (ptype::image_body_hash(p + sizeof(*h), bytes - sizeof(*h)) != h->body_checksum)
        return PTYPE_E_BAD_IMAGE;
    return PTYPE_OK;
}

static const ptype_service* find_service(const void* image, const char* service) {
    const ptype_image_header* h = static_cast<const ptype_image_header*>(image);
    if (!h->host_sections) return nullptr;
    const unsigned char* p = static_cast<const unsigned char*>(image);
    const ptype_service* sv = reinterpret_cast<const ptype_service*>(p + h->off_services);
    const char* pool = reinterpret_cast<const char*>(p + h->off_strings);
    uint64_t key = ptype_b200_service_key(service, h->key_salt);
    const ptype_service* end = sv + h->n_services;
    const ptype_service* it = std::lower_bound(
        sv, end, key, [](const ptype_service& s, uint64_t k) { return s.key < k; });
    if (it == end || it->key != key || std::strcmp(pool + it->name_off, service)) return nullptr;
    return it;
}

// What getSelectedNodes() (cluster/rpc.go:293-297) holds after handleNewNodes
// applied selectNodes (rpc.go:226-236,252-264) to this snapshot's list.
int ptype_b200_image_selected(const void* image, const char* service, uint32_t* selected,
                              size_t cap, uint32_t* k_out) {
    if (!image || !service || !k_out) return PTYPE_E_INVAL;
    const ptype_image_header* h = static_cast<const ptype_image_header*>(image);
    const ptype_service* s = find_service(image, service);
    if (!s) return PTYPE_E_NOT_FOUND;
    const unsigned char* p = static_cast<const unsigned char*>(image);
    const uint32_t* conn = reinterpret_cast<const uint32_t*>(p + h->off_conn_hash);
    const uint32_t n = s->n_nodes, m = h->max_connections;
    const bool identity = m == 0 || n <= m;
    const uint32_t k = identity ? n : m;
    for (uint32_t i = 0; i < k && i < cap; ++i)
        selected[i] = s->node_base + (identity ? i : conn[i] % n);
    *k_out = k;
    return PTYPE_OK;
}

int ptype_b200_image_node(const void* image, uint32_t node_id, char* address, size_t address_cap,
                          int* port, char* service, size_t service_cap, char* node_name,
                          size_t node_name_cap) {
    if (!image) return PTYPE_E_INVAL;
    const ptype_image_header* h = static_cast<const ptype_image_header*>(image);
    if (node_id >= h->n_nodes || !h->host_sections) return PTYPE_E_NOT_FOUND;
    const unsigned char* p = static_cast<const unsigned char*>(image);
    const ptype_node* nd = reinterpret_cast<const ptype_node*>(p + h->off_nodes) + node_id;
    const ptype_service* sv = reinterpret_cast<const ptype_service*>(p + h->off_services);
    const char* pool = reinterpret_cast<const char*>(p + h->off_strings);
    if (address && address_cap) std::snprintf(address, address_cap, "%s", pool + nd->address_off);
    if (port) *port = nd->port;
    if (service && service_cap)
        std::snprintf(service, service_cap, "%s", pool + sv[nd->service].name_off);
    if (node_name && node_name_cap) std::snprintf(node_name, node_name_cap, "%s", pool + nd->name_off);
    return PTYPE_OK;
}

}  // extern "C"
