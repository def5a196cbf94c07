/*
 * ptype_oracle.c — see ptype_oracle.h.  TEST INFRASTRUCTURE ONLY (checker and
 * timed CPU baseline); the product never links this file.
 *
 * Plain C11, no dependencies beyond libc + pthreads.  Integer arithmetic that
 * Go defines and C leaves undefined (signed overflow, INT64_MIN % -1) is done
 * in uint64_t or special-cased.
 */
#define _POSIX_C_SOURCE 200809L
#include "ptype_oracle.h"

#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

/* ======================================================== Go stdlib pieces */

/* hash/fnv New32a: offset 0x811c9dc5, prime 0x01000193, xor then multiply.
 * Call site: cluster/rpc.go:267-269. */
uint32_t po_fnv1a32(const void* p, size_t n) {
    const unsigned char* s = (const unsigned char*)p;
    uint32_t h = 0x811c9dc5u;
    for (size_t i = 0; i < n; i++) {
        h ^= s[i];
        h *= 0x01000193u;
    }
    return h;
}

/* hash/fnv New64a with the offset basis xor-ed by a salt (salt 0 = Go's New64a).
 * Not in the reference: it is the registry-key hash of the build. */
uint64_t po_fnv1a64_salted(const void* p, size_t n, uint64_t salt) {
    const unsigned char* s = (const unsigned char*)p;
    uint64_t h = 0xcbf29ce484222325ull ^ salt;
    for (size_t i = 0; i < n; i++) {
        h ^= s[i];
        h *= 0x100000001b3ull;
    }
    return h;
}

/* strconv.Itoa (cluster/rpc.go:268). Returns length, NUL-terminates. */
int po_itoa(int64_t v, char* out) {
    char tmp[24];
    int n = 0;
    uint64_t u = v < 0 ? (uint64_t)0 - (uint64_t)v : (uint64_t)v;
    do {
        tmp[n++] = (char)('0' + u % 10);
        u /= 10;
    } while (u);
    int len = 0;
    if (v < 0) out[len++] = '-';
    while (n) out[len++] = tmp[--n];
    out[len] = 0;
    return len;
}

/* path/filepath.Clean on a '/'-separated path (lexical only). */
static int po_path_clean(const char* in, char* out, size_t cap) {
    size_t n = strlen(in);
    if (n == 0) {
        if (cap < 2) return -1;
        strcpy(out, ".");
        return 1;
    }
    int rooted = in[0] == '/';
    size_t w = 0, r = 0, dotdot = 0;
    if (cap < n + 2) return -1;
    if (rooted) {
        out[w++] = '/';
        r = 1;
        dotdot = 1;
    }
    while (r < n) {
        if (in[r] == '/') {
            r++;
        } else if (in[r] == '.' && (r + 1 == n || in[r + 1] == '/')) {
            r++;
        } else if (in[r] == '.' && in[r + 1] == '.' && (r + 2 == n || in[r + 2] == '/')) {
            r += 2;
            if (w > dotdot) {
                w--;
                while (w > dotdot && out[w] != '/') w--;
            } else if (!rooted) {
                if (w > 0) out[w++] = '/';
                out[w++] = '.';
                out[w++] = '.';
                dotdot = w;
            }
        } else {
            if ((rooted && w != 1) || (!rooted && w != 0)) out[w++] = '/';
            while (r < n && in[r] != '/') out[w++] = in[r++];
        }
    }
    if (w == 0) out[w++] = '.';
    out[w] = 0;
    return (int)w;
}

/* etcdKey (cluster/registry.go:168-170): filepath.Join(elems...) + "/".
 * Join drops empty elements, joins with '/', then Cleans. */
int po_etcd_key(char* out, size_t cap, const char* const* elems, int n) {
    size_t total = 0;
    for (int i = 0; i < n; i++) total += strlen(elems[i]) + 1;
    char* joined = (char*)malloc(total + 1);
    if (!joined) return -1;
    size_t w = 0;
    int any = 0;
    for (int i = 0; i < n; i++) {
        if (elems[i][0] == 0) continue;
        if (any) joined[w++] = '/';
        size_t l = strlen(elems[i]);
        memcpy(joined + w, elems[i], l);
        w += l;
        any = 1;
    }
    joined[w] = 0;
    int len;
    if (!any) {
        if (cap < 2) {
            free(joined);
            return -1;
        }
        out[0] = 0;
        len = 0;
    } else {
        len = po_path_clean(joined, out, cap);
    }
    free(joined);
    if (len < 0 || (size_t)len + 2 > cap) return -1;
    out[len++] = '/';
    out[len] = 0;
    return len;
}

/* ============================================================= ordered KV */

struct po_kv {
    char** keys;
    char** vals;
    size_t n, cap;
};

po_kv* po_kv_new(void) { return (po_kv*)calloc(1, sizeof(po_kv)); }

void po_kv_free(po_kv* kv) {
    if (!kv) return;
    for (size_t i = 0; i < kv->n; i++) {
        free(kv->keys[i]);
        free(kv->vals[i]);
    }
    free(kv->keys);
    free(kv->vals);
    free(kv);
}

/* first index with key >= k (byte order, like etcd's key sort) */
static size_t po_kv_lower(const po_kv* kv, const char* k) {
    size_t lo = 0, hi = kv->n;
    while (lo < hi) {
        size_t mid = (lo + hi) / 2;
        if (strcmp(kv->keys[mid], k) < 0)
            lo = mid + 1;
        else
            hi = mid;
    }
    return lo;
}

void po_kv_put(po_kv* kv, const char* key, const char* value) {
    size_t i = po_kv_lower(kv, key);
    if (i < kv->n && strcmp(kv->keys[i], key) == 0) {
        free(kv->vals[i]);
        kv->vals[i] = strdup(value);
        return;
    }
    if (kv->n == kv->cap) {
        kv->cap = kv->cap ? kv->cap * 2 : 64;
        kv->keys = (char**)realloc(kv->keys, kv->cap * sizeof(char*));
        kv->vals = (char**)realloc(kv->vals, kv->cap * sizeof(char*));
    }
    memmove(kv->keys + i + 1, kv->keys + i, (kv->n - i) * sizeof(char*));
    memmove(kv->vals + i + 1, kv->vals + i, (kv->n - i) * sizeof(char*));
    kv->keys[i] = strdup(key);
    kv->vals[i] = strdup(value);
    kv->n++;
}

int po_kv_delete(po_kv* kv, const char* key) {
    size_t i = po_kv_lower(kv, key);
    if (i >= kv->n || strcmp(kv->keys[i], key) != 0) return 0;
    free(kv->keys[i]);
    free(kv->vals[i]);
    memmove(kv->keys + i, kv->keys + i + 1, (kv->n - i - 1) * sizeof(char*));
    memmove(kv->vals + i, kv->vals + i + 1, (kv->n - i - 1) * sizeof(char*));
    kv->n--;
    return 1;
}

/* clientv3.WithPrefix + SortByKey/SortAscend (cluster/registry.go:88-91). */
size_t po_kv_range_prefix(const po_kv* kv, const char* prefix, size_t* first) {
    size_t plen = strlen(prefix);
    size_t i = po_kv_lower(kv, prefix);
    *first = i;
    size_t c = 0;
    while (i + c < kv->n && strncmp(kv->keys[i + c], prefix, plen) == 0) c++;
    return c;
}
const char* po_kv_key_at(const po_kv* kv, size_t i) { return kv->keys[i]; }
const char* po_kv_value_at(const po_kv* kv, size_t i) { return kv->vals[i]; }
size_t po_kv_len(const po_kv* kv) { return kv->n; }

/* ================================================================ Node JSON */

/* json.Marshal(Node{Address,Port}) (cluster/registry.go:52-53):
 * {"address":"...","port":N}; Go escapes ", \, control bytes and <,>,&. */
int po_node_marshal(const po_node* nd, char* out, size_t cap) {
    size_t w = 0;
#define PUT(c)                        \
    do {                              \
        if (w + 1 >= cap) return -1;  \
        out[w++] = (char)(c);         \
    } while (0)
#define PUTS(s)                                   \
    do {                                          \
        for (const char* q_ = (s); *q_; q_++) PUT(*q_); \
    } while (0)
    PUTS("{\"address\":\"");
    for (const unsigned char* s = (const unsigned char*)nd->address; *s; s++) {
        char buf[8];
        if (*s == '"' || *s == '\\') {
            PUT('\\');
            PUT(*s);
        } else if (*s == '\n') {
            PUTS("\\n");
        } else if (*s == '\r') {
            PUTS("\\r");
        } else if (*s == '\t') {
            PUTS("\\t");
        } else if (*s < 0x20 || *s == '<' || *s == '>' || *s == '&') {
            snprintf(buf, sizeof buf, "\\u%04x", *s);
            PUTS(buf);
        } else {
            PUT(*s);
        }
    }
    PUTS("\",\"port\":");
    char num[24];
    po_itoa(nd->port, num);
    PUTS(num);
    PUT('}');
    out[w] = 0;
    return (int)w;
#undef PUT
#undef PUTS
}

typedef struct {
    const char* p;
} jcur;
static void j_ws(jcur* c) {
    while (*c->p == ' ' || *c->p == '\t' || *c->p == '\n' || *c->p == '\r') c->p++;
}
static int j_hex4(const char* p, unsigned* out) {
    unsigned v = 0;
    for (int i = 0; i < 4; i++) {
        char ch = p[i];
        v <<= 4;
        if (ch >= '0' && ch <= '9')
            v |= (unsigned)(ch - '0');
        else if (ch >= 'a' && ch <= 'f')
            v |= (unsigned)(ch - 'a' + 10);
        else if (ch >= 'A' && ch <= 'F')
            v |= (unsigned)(ch - 'A' + 10);
        else
            return -1;
    }
    *out = v;
    return 0;
}
/* parse a JSON string at c->p ('"' first); writes up to cap-1 bytes when out != NULL */
static int j_string(jcur* c, char* out, size_t cap) {
    if (*c->p != '"') return -1;
    c->p++;
    size_t w = 0;
#define EMIT(b)                                  \
    do {                                         \
        if (out) {                               \
            if (w + 1 >= cap) return -1;         \
            out[w] = (char)(b);                  \
        }                                        \
        w++;                                     \
    } while (0)
    while (*c->p && *c->p != '"') {
        unsigned char ch = (unsigned char)*c->p++;
        if (ch < 0x20) return -1;
        if (ch != '\\') {
            EMIT(ch);
            continue;
        }
        char e = *c->p++;
        switch (e) {
            case '"': EMIT('"'); break;
            case '\\': EMIT('\\'); break;
            case '/': EMIT('/'); break;
            case 'b': EMIT('\b'); break;
            case 'f': EMIT('\f'); break;
            case 'n': EMIT('\n'); break;
            case 'r': EMIT('\r'); break;
            case 't': EMIT('\t'); break;
            case 'u': {
                unsigned cp;
                if (j_hex4(c->p, &cp)) return -1;
                c->p += 4;
                if (cp >= 0xD800 && cp < 0xDC00 && c->p[0] == '\\' && c->p[1] == 'u') {
                    unsigned lo;
                    if (j_hex4(c->p + 2, &lo) == 0 && lo >= 0xDC00 && lo < 0xE000) {
                        cp = 0x10000 + ((cp - 0xD800) << 10) + (lo - 0xDC00);
                        c->p += 6;
                    }
                }
                if (cp >= 0xD800 && cp < 0xE000) cp = 0xFFFD;
                if (cp < 0x80) {
                    EMIT(cp);
                } else if (cp < 0x800) {
                    EMIT(0xC0 | (cp >> 6));
                    EMIT(0x80 | (cp & 0x3F));
                } else if (cp < 0x10000) {
                    EMIT(0xE0 | (cp >> 12));
                    EMIT(0x80 | ((cp >> 6) & 0x3F));
                    EMIT(0x80 | (cp & 0x3F));
                } else {
                    EMIT(0xF0 | (cp >> 18));
                    EMIT(0x80 | ((cp >> 12) & 0x3F));
                    EMIT(0x80 | ((cp >> 6) & 0x3F));
                    EMIT(0x80 | (cp & 0x3F));
                }
                break;
            }
            default: return -1;
        }
    }
#undef EMIT
    if (*c->p != '"') return -1;
    c->p++;
    if (out) out[w] = 0;
    return 0;
}
static int j_skip_value(jcur* c);
static int j_skip_number(jcur* c, int* is_int) {
    const char* s = c->p;
    *is_int = 1;
    if (*c->p == '-') c->p++;
    if (*c->p < '0' || *c->p > '9') return -1;
    if (*c->p == '0') {
        c->p++;
    } else {
        while (*c->p >= '0' && *c->p <= '9') c->p++;
    }
    if (*c->p == '.') {
        *is_int = 0;
        c->p++;
        if (*c->p < '0' || *c->p > '9') return -1;
        while (*c->p >= '0' && *c->p <= '9') c->p++;
    }
    if (*c->p == 'e' || *c->p == 'E') {
        *is_int = 0;
        c->p++;
        if (*c->p == '+' || *c->p == '-') c->p++;
        if (*c->p < '0' || *c->p > '9') return -1;
        while (*c->p >= '0' && *c->p <= '9') c->p++;
    }
    return c->p > s ? 0 : -1;
}
static int j_skip_value(jcur* c) {
    j_ws(c);
    if (*c->p == '"') return j_string(c, NULL, 0);
    if (*c->p == '{' || *c->p == '[') {
        char open = *c->p, close = open == '{' ? '}' : ']';
        c->p++;
        j_ws(c);
        if (*c->p == close) {
            c->p++;
            return 0;
        }
        for (;;) {
            j_ws(c);
            if (open == '{') {
                if (j_string(c, NULL, 0)) return -1;
                j_ws(c);
                if (*c->p++ != ':') return -1;
            }
            if (j_skip_value(c)) return -1;
            j_ws(c);
            if (*c->p == ',') {
                c->p++;
                continue;
            }
            if (*c->p == close) {
                c->p++;
                return 0;
            }
            return -1;
        }
    }
    if (!strncmp(c->p, "true", 4)) {
        c->p += 4;
        return 0;
    }
    if (!strncmp(c->p, "false", 5)) {
        c->p += 5;
        return 0;
    }
    if (!strncmp(c->p, "null", 4)) {
        c->p += 4;
        return 0;
    }
    int is_int;
    return j_skip_number(c, &is_int);
}
static int j_key_is(const char* key, const char* want) { /* json field match folds case */
    for (; *key && *want; key++, want++) {
        char a = *key, b = *want;
        if (a >= 'A' && a <= 'Z') a = (char)(a + 32);
        if (a != b) return 0;
    }
    return *key == 0 && *want == 0;
}

/* json.Unmarshal(value, &node) (cluster/registry.go:102,161): only the fields
 * present are assigned; `null` leaves a field as is; a non-integer port or a
 * non-string address is an error. */
int po_node_unmarshal(const char* json, po_node* nd) {
    jcur c = {json};
    j_ws(&c);
    if (!strncmp(c.p, "null", 4)) {
        c.p += 4;
        j_ws(&c);
        return *c.p ? -1 : 0;
    }
    if (*c.p++ != '{') return -1;
    j_ws(&c);
    if (*c.p == '}') {
        c.p++;
        j_ws(&c);
        return *c.p ? -1 : 0;
    }
    for (;;) {
        char key[64];
        j_ws(&c);
        if (j_string(&c, key, sizeof key)) {
            /* over-long keys cannot name a Node field: skip them */
            return -1;
        }
        j_ws(&c);
        if (*c.p++ != ':') return -1;
        j_ws(&c);
        if (j_key_is(key, "address")) {
            if (!strncmp(c.p, "null", 4)) {
                c.p += 4;
            } else if (*c.p == '"') {
                char tmp[sizeof nd->address];
                if (j_string(&c, tmp, sizeof tmp)) return -1;
                strcpy(nd->address, tmp);
            } else {
                return -1;
            }
        } else if (j_key_is(key, "port")) {
            if (!strncmp(c.p, "null", 4)) {
                c.p += 4;
            } else {
                const char* s = c.p;
                int is_int;
                if (j_skip_number(&c, &is_int) || !is_int) return -1;
                char num[32];
                size_t l = (size_t)(c.p - s);
                if (l >= sizeof num) return -1;
                memcpy(num, s, l);
                num[l] = 0;
                nd->port = strtoll(num, NULL, 10);
            }
        } else if (j_skip_value(&c)) {
            return -1;
        }
        j_ws(&c);
        if (*c.p == ',') {
            c.p++;
            continue;
        }
        if (*c.p == '}') {
            c.p++;
            break;
        }
        return -1;
    }
    j_ws(&c);
    return *c.p ? -1 : 0;
}

/* ================================================================= registry */

static const char* const PO_SERVICES_PREFIX = "services"; /* registry.go:15 */

/* Register (cluster/registry.go:51-86): key services/<svc>/<node>/ , JSON value.
 * The lease and keep-alive (registry.go:59-83) have no effect on the stored pair. */
int po_register(po_kv* kv, const char* service, const char* node_name, const char* host,
                int64_t port) {
    po_node nd;
    memset(&nd, 0, sizeof nd);
    if (strlen(host) >= sizeof nd.address) return -1;
    strcpy(nd.address, host);
    nd.port = port;
    char val[1024];
    if (po_node_marshal(&nd, val, sizeof val) < 0) return -1;
    const char* elems[3] = {PO_SERVICES_PREFIX, service, node_name};
    char key[1024];
    if (po_etcd_key(key, sizeof key, elems, 3) < 0) return -1;
    po_kv_put(kv, key, val);
    return 0;
}

/* Services (cluster/registry.go:93-117): ONE range over the prefix "services"
 * (no trailing slash), ONE reused Node (line 99), service = Split(key,"/")[1]. */
long po_services(const po_kv* kv, po_service_list** out) {
    size_t first;
    size_t cnt = po_kv_range_prefix(kv, PO_SERVICES_PREFIX, &first);
    po_service_list* lists = NULL;
    size_t nl = 0, capl = 0;
    po_node node;
    memset(&node, 0, sizeof node);
    for (size_t i = 0; i < cnt; i++) {
        if (po_node_unmarshal(po_kv_value_at(kv, first + i), &node)) {
            po_services_free(lists, nl);
            return -1;
        }
        const char* key = po_kv_key_at(kv, first + i);
        const char* s1 = strchr(key, '/');
        if (!s1) continue; /* len(Split) <= 1 */
        s1++;
        const char* s2 = strchr(s1, '/');
        size_t sl = s2 ? (size_t)(s2 - s1) : strlen(s1);
        char name[256];
        if (sl >= sizeof name) sl = sizeof name - 1;
        memcpy(name, s1, sl);
        name[sl] = 0;
        size_t j = 0;
        for (; j < nl; j++)
            if (!strcmp(lists[j].name, name)) break;
        if (j == nl) {
            if (nl == capl) {
                capl = capl ? capl * 2 : 16;
                lists = (po_service_list*)realloc(lists, capl * sizeof *lists);
            }
            memset(&lists[nl], 0, sizeof lists[nl]);
            strcpy(lists[nl].name, name);
            nl++;
        }
        po_service_list* L = &lists[j];
        if ((L->n & (L->n + 1)) == 0 || L->n == 0) { /* grow at 0,1,3,7,... */
            size_t ncap = L->n ? (L->n + 1) * 2 : 1;
            L->nodes = (po_node*)realloc(L->nodes, ncap * sizeof(po_node));
        }
        L->nodes[L->n++] = node;
    }
    *out = lists;
    return (long)nl;
}

void po_services_free(po_service_list* l, size_t n) {
    if (!l) return;
    for (size_t i = 0; i < n; i++) free(l[i].nodes);
    free(l);
}

/* nodes (cluster/registry.go:152-166): range services/<svc>/ — the trailing
 * slash is what keeps "foo" from matching "foo_broken_prefix_case"
 * (cluster/registry_test.go:209-236). Fresh Node per element (line 159). */
long po_nodes(const po_kv* kv, const char* service, po_node** out) {
    const char* elems[2] = {PO_SERVICES_PREFIX, service};
    char key[1024];
    if (po_etcd_key(key, sizeof key, elems, 2) < 0) return -1;
    size_t first;
    size_t cnt = po_kv_range_prefix(kv, key, &first);
    po_node* nodes = (po_node*)calloc(cnt ? cnt : 1, sizeof(po_node));
    for (size_t i = 0; i < cnt; i++) {
        if (po_node_unmarshal(po_kv_value_at(kv, first + i), &nodes[i])) {
            free(nodes);
            return -1;
        }
    }
    *out = nodes;
    return (long)cnt;
}

/* ================================================================= balancer */

/* hashConnectionIndex (cluster/rpc.go:266-270):
 *   h := fnv.New32a(); h.Write([]byte(localAddr + strconv.Itoa(connNumber)))
 *   return int(h.Sum32()) % nodeSize            (int is 64-bit: never negative) */
int64_t po_hash_connection_index(const char* local_addr, int64_t conn_number, int64_t node_size) {
    char buf[512];
    size_t l = strlen(local_addr);
    if (l > sizeof buf - 24) l = sizeof buf - 24;
    memcpy(buf, local_addr, l);
    int dl = po_itoa(conn_number, buf + l);
    uint32_t h = po_fnv1a32(buf, l + (size_t)dl);
    return (int64_t)h % node_size;
}

/* selectNodes (cluster/rpc.go:252-264). */
long po_select_nodes(const char* local_addr, size_t n_nodes, int64_t max_connections,
                     uint32_t* selected_idx, size_t cap) {
    if (max_connections < 0) return -1;
    if ((int64_t)n_nodes <= max_connections || max_connections == 0) { /* rpc.go:253-255 */
        if (selected_idx)
            for (size_t i = 0; i < n_nodes && i < cap; i++) selected_idx[i] = (uint32_t)i;
        return (long)n_nodes;
    }
    size_t len = 0;
    for (int64_t i = 0; (int64_t)len < max_connections; i++) { /* rpc.go:258-261: no dedup */
        int64_t index = po_hash_connection_index(local_addr, i, (int64_t)n_nodes);
        if (selected_idx && len < cap) selected_idx[len] = (uint32_t)index;
        len++;
    }
    return (long)len;
}

/* roundRobinSelect (cluster/rpc.go:176-183): index is the value AddUint64
 * returned; clients[int(index)%len(clients)].  int(index) reinterprets the bits. */
int64_t po_round_robin(uint64_t seq_after_add, size_t n_clients) {
    if (n_clients == 0) return -1; /* rpc.go:178-180 → nil → ErrNoClientAvailable */
    int64_t idx = (int64_t)seq_after_add;
    int64_t n = (int64_t)n_clients;
    int64_t r = idx % n; /* C99 and Go both truncate toward zero; n > 0 so no UB */
    if (r < 0) return -2; /* Go: index out of range panic */
    return r;
}

/* ============================================================ method bodies */

/* Calculator.Multiply (example/calculator/calculator.go:9-12): Go int wraps. */
int64_t po_multiply(int64_t a, int64_t b) { return (int64_t)((uint64_t)a * (uint64_t)b); }
/* Calculator.Add — defined by the build, same wrapping rule. */
int64_t po_add(int64_t a, int64_t b) { return (int64_t)((uint64_t)a + (uint64_t)b); }

static int64_t po_min(int64_t a, int64_t b) { return a < b ? a : b; } /* prime.go:27-32 */

/* Go's %: truncated, and MinInt64 % -1 == 0 without trapping. */
static int64_t po_go_rem(int64_t a, int64_t b) {
    if (b == -1) return 0;
    return a % b;
}

/* Prime.Check (example/optimus/prime.go:15-25), time.Sleep dropped (D4). */
int64_t po_check(int64_t min, int64_t max, int64_t target) {
    int64_t bound = po_min(max, target);
    for (int64_t i = min; i < bound; i++) {
        if (i != 0 && po_go_rem(target, i) == 0) return i;
    }
    return target;
}

uint64_t po_check_width(int64_t min, int64_t max, int64_t target) {
    int64_t bound = po_min(max, target);
    if (min >= bound) return 0;
    return (uint64_t)bound - (uint64_t)min;
}

/* splitWork (example/optimus/coordinator/coordinator.go:67-73):
 *   min := 2; for i := 10; i < target+10; i += 10 { checkRange(min,i,target); min = i } */
size_t po_split_work(int64_t target, int64_t (*out)[3], size_t cap) {
    size_t c = 0;
    int64_t min = 2;
    int64_t limit = (int64_t)((uint64_t)target + 10u); /* Go wraps */
    for (int64_t i = 10; i < limit; i += 10) {
        if (out && c < cap) {
            out[c][0] = min;
            out[c][1] = i;
            out[c][2] = target;
        }
        c++;
        min = i;
        if (i > INT64_MAX - 10) break; /* i += 10 would wrap negative and keep looping in Go; unreachable for sane targets */
    }
    return c;
}

/* ================================================================= snapshot */

#define PO_RECV_CALCULATOR 1u
#define PO_RECV_PRIME 2u

typedef struct {
    char name[256];
    uint64_t key;
    uint32_t node_base; /* global id of nodes[0] */
    uint32_t n_nodes;
    uint32_t* selected; /* global node ids: getSelectedNodes() */
    uint32_t k;
    uint32_t receivers; /* PO_RECV_* bits */
    uint32_t created;   /* position before the sort by key (build-time bookkeeping) */
} po_client;

#define PO_MAX_METHODS 8
struct po_snapshot {
    uint32_t method_hash[PO_MAX_METHODS]; /* FNV-1a-32("Recv.Method"), hashed once per snapshot */
    uint32_t recv_hash[PO_MAX_METHODS];   /* FNV-1a-32("Recv") */
    po_client* clients; /* sorted by key */
    size_t n_clients;
    uint32_t* index;    /* open-addressing hash over clients[].key: the Go side of this lookup is
                         * a map from service name to *Client, i.e. a hash table */
    size_t index_mask;
    po_node* nodes; /* global node array, in `services` range order */
    uint32_t* node_service;
    size_t n_nodes;
    uint64_t salt;
};

static void po_snapshot_hash_methods(po_snapshot* s);

static int po_client_cmp(const void* a, const void* b) {
    uint64_t ka = ((const po_client*)a)->key, kb = ((const po_client*)b)->key;
    return ka < kb ? -1 : ka > kb;
}

/* Host-side state after: services := Registry.Services() (registry.go:93-117);
 * for each service: NewClient(service, cfg) → handleNewNodes(nodes) →
 * selectNodes (rpc.go:226-236,252-264).  Node lists per service are what
 * nodes(service) returns (registry.go:152-166); global ids number the
 * services/ range in key order. */
po_snapshot* po_snapshot_build(const po_kv* kv, const char* local_addr, int64_t max_connections,
                               const char* const* declared, size_t n_declared, uint64_t salt) {
    if (max_connections < 0) return NULL;
    po_snapshot* s = (po_snapshot*)calloc(1, sizeof *s);
    s->salt = salt;
    /* global node numbering: every key under "services/" in order */
    size_t first;
    size_t cnt = po_kv_range_prefix(kv, "services/", &first);
    s->nodes = (po_node*)calloc(cnt ? cnt : 1, sizeof(po_node));
    s->node_service = (uint32_t*)calloc(cnt ? cnt : 1, sizeof(uint32_t));
    s->n_nodes = cnt;
    size_t capc = 16;
    s->clients = (po_client*)calloc(capc, sizeof(po_client));
    for (size_t i = 0; i < cnt; i++) {
        const char* key = po_kv_key_at(kv, first + i);
        const char* s1 = key + strlen("services/");
        const char* s2 = strchr(s1, '/');
        size_t sl = s2 ? (size_t)(s2 - s1) : strlen(s1);
        char name[256];
        if (sl >= sizeof name) sl = sizeof name - 1;
        memcpy(name, s1, sl);
        name[sl] = 0;
        if (s->n_clients == 0 || strcmp(s->clients[s->n_clients - 1].name, name)) {
            if (s->n_clients == capc) {
                capc *= 2;
                s->clients = (po_client*)realloc(s->clients, capc * sizeof(po_client));
            }
            po_client* c = &s->clients[s->n_clients++];
            memset(c, 0, sizeof *c);
            strcpy(c->name, name);
            c->node_base = (uint32_t)i;
        }
        po_client* c = &s->clients[s->n_clients - 1];
        c->n_nodes++;
        if (po_node_unmarshal(po_kv_value_at(kv, first + i), &s->nodes[i])) {
            po_snapshot_free(s);
            return NULL;
        }
        s->node_service[i] = (uint32_t)(s->n_clients - 1);
    }
    for (size_t d = 0; d < n_declared; d++) {
        size_t j = 0;
        for (; j < s->n_clients; j++)
            if (!strcmp(s->clients[j].name, declared[d])) break;
        if (j < s->n_clients) continue;
        if (s->n_clients == capc) {
            capc *= 2;
            s->clients = (po_client*)realloc(s->clients, capc * sizeof(po_client));
        }
        po_client* c = &s->clients[s->n_clients++];
        memset(c, 0, sizeof *c);
        snprintf(c->name, sizeof c->name, "%s", declared[d]);
    }
    for (size_t j = 0; j < s->n_clients; j++) {
        po_client* c = &s->clients[j];
        c->key = po_fnv1a64_salted(c->name, strlen(c->name), salt);
        long k = po_select_nodes(local_addr, c->n_nodes, max_connections, NULL, 0);
        c->k = (uint32_t)k;
        c->selected = (uint32_t*)calloc(k ? (size_t)k : 1, sizeof(uint32_t));
        po_select_nodes(local_addr, c->n_nodes, max_connections, c->selected, (size_t)k);
        for (long i = 0; i < k; i++) c->selected[i] += c->node_base;
    }
    /* node_service indexes clients by creation order; renumber after the sort by key */
    for (size_t j = 0; j < s->n_clients; j++) s->clients[j].created = (uint32_t)j;
    qsort(s->clients, s->n_clients, sizeof(po_client), po_client_cmp);
    uint32_t* now_at = (uint32_t*)calloc(s->n_clients ? s->n_clients : 1, sizeof(uint32_t));
    for (size_t j = 0; j < s->n_clients; j++) now_at[s->clients[j].created] = (uint32_t)j;
    for (size_t i = 0; i < s->n_nodes; i++) s->node_service[i] = now_at[s->node_service[i]];
    free(now_at);
    /* hash index, at most half full, linear probing; entries are client positions + 1 */
    size_t cap = 16;
    while (cap < 2 * s->n_clients) cap <<= 1;
    s->index = (uint32_t*)calloc(cap, sizeof(uint32_t));
    s->index_mask = cap - 1;
    for (size_t j = 0; j < s->n_clients; j++) {
        size_t h = (size_t)((s->clients[j].key * 0x9E3779B97F4A7C15ull) >> 20) & s->index_mask;
        while (s->index[h]) h = (h + 1) & s->index_mask;
        s->index[h] = (uint32_t)j + 1;
    }
    po_snapshot_hash_methods(s);
    return s;
}

void po_snapshot_free(po_snapshot* s) {
    if (!s) return;
    for (size_t j = 0; j < s->n_clients; j++) free(s->clients[j].selected);
    free(s->clients);
    free(s->index);
    free(s->nodes);
    free(s->node_service);
    free(s);
}

static po_client* po_find_name(const po_snapshot* s, const char* service) {
    for (size_t j = 0; j < s->n_clients; j++)
        if (!strcmp(s->clients[j].name, service)) return &s->clients[j];
    return NULL;
}

static const po_client* po_find_key(const po_snapshot* s, uint64_t key) {
    size_t h = (size_t)((key * 0x9E3779B97F4A7C15ull) >> 20) & s->index_mask;
    for (uint32_t e; (e = s->index[h]) != 0; h = (h + 1) & s->index_mask)
        if (s->clients[e - 1].key == key) return &s->clients[e - 1];
    return NULL;
}

/* rpc.Register(new(calculator.Calculator)) (example/calculator/server/server.go:16-17),
 * rpc.Register(new(prime.Prime)) (example/optimus/worker/worker.go:16-17). */
int po_snapshot_expose(po_snapshot* s, const char* service, const char* receiver) {
    po_client* c = po_find_name(s, service);
    if (!c) return -1;
    if (!strcmp(receiver, "Calculator"))
        c->receivers |= PO_RECV_CALCULATOR;
    else if (!strcmp(receiver, "Prime"))
        c->receivers |= PO_RECV_PRIME;
    else
        return -1;
    return 0;
}

size_t po_snapshot_n_nodes(const po_snapshot* s) { return s->n_nodes; }

long po_snapshot_selected(const po_snapshot* s, const char* service, uint32_t* out, size_t cap) {
    const po_client* c = po_find_name(s, service);
    if (!c) return -1;
    for (uint32_t i = 0; i < c->k && i < cap; i++) out[i] = c->selected[i];
    return (long)c->k;
}

int po_snapshot_node(const po_snapshot* s, uint32_t node_id, po_node* out, char* service,
                     size_t cap) {
    if (node_id >= s->n_nodes) return -1;
    *out = s->nodes[node_id];
    if (service) snprintf(service, cap, "%s", s->clients[s->node_service[node_id]].name);
    return 0;
}

/* ------------------------------------------------------------ one call */

typedef struct {
    const char* recv;
    uint32_t recv_bit;
    const char* method;
    int opcode;
} po_method_def;

/* The exported methods of the two example receivers (net/rpc registers every
 * exported method with the (args, *reply) error shape): calculator.go:9,
 * prime.go:15, plus the build-defined Add (D1). */
static const po_method_def PO_METHODS[] = {
    {"Calculator", PO_RECV_CALCULATOR, "Multiply", PTYPE_OP_MULTIPLY},
    {"Calculator", PO_RECV_CALCULATOR, "Add", PTYPE_OP_ADD},
    {"Prime", PO_RECV_PRIME, "Check", PTYPE_OP_CHECK},
};
#define PO_N_METHODS (sizeof PO_METHODS / sizeof PO_METHODS[0])

/* What rpc.Register builds once per server: the method table, here as name hashes. */
static void po_snapshot_hash_methods(po_snapshot* s) {
    for (size_t m = 0; m < PO_N_METHODS; m++) {
        const po_method_def* d = &PO_METHODS[m];
        char full[128];
        int l = snprintf(full, sizeof full, "%s.%s", d->recv, d->method);
        s->method_hash[m] = po_fnv1a32(full, (size_t)l);
        s->recv_hash[m] = po_fnv1a32(d->recv, strlen(d->recv));
    }
}

void po_method_key(const char* service_method, uint32_t* method, uint32_t* recv) {
    *method = po_fnv1a32(service_method, strlen(service_method));
    const char* dot = strrchr(service_method, '.'); /* strings.LastIndex(sm, ".") [net/rpc] */
    *recv = dot ? po_fnv1a32(service_method, (size_t)(dot - service_method)) : PTYPE_RECV_ILL_FORMED;
}

static void po_execute(int opcode, const int64_t a[3], uint64_t check_limit, ptype_reply* r) {
    switch (opcode) {
        case PTYPE_OP_MULTIPLY: r->value = po_multiply(a[0], a[1]); break;
        case PTYPE_OP_ADD: r->value = po_add(a[0], a[1]); break;
        case PTYPE_OP_CHECK:
            if (po_check_width(a[0], a[1], a[2]) > check_limit) {
                r->status = PTYPE_ST_RANGE_LIMIT;
                return;
            }
            r->value = po_check(a[0], a[1], a[2]);
            break;
    }
}

/* Client.Call, one attempt (cluster/rpc.go:59-67), then what the net/rpc server
 * at the chosen node does with the request header [Go 1.13 net/rpc
 * server.readRequestHeader]: LastIndex(".") < 0 → ill-formed; unknown receiver →
 * can't find service; unknown method → can't find method. */
static void po_call_client(const po_snapshot* s, const po_client* c, uint64_t seq, uint32_t method,
                           uint32_t recv, const int64_t a[3], uint64_t check_limit,
                           ptype_reply* r) {
    r->value = 0;
    r->node = PTYPE_NO_NODE;
    if (!c) {
        r->status = PTYPE_ST_NO_SERVICE;
        return;
    }
    int64_t idx = po_round_robin(seq, c->k);
    if (idx == -1) {
        r->status = PTYPE_ST_NO_CLIENT;
        return;
    }
    if (idx == -2) {
        r->status = PTYPE_ST_BAD_INDEX;
        return;
    }
    r->node = c->selected[idx];
    if (recv == PTYPE_RECV_ILL_FORMED) {
        r->status = PTYPE_ST_ILL_FORMED;
        return;
    }
    int recv_known = 0;
    for (size_t m = 0; m < PO_N_METHODS; m++) {
        const po_method_def* d = &PO_METHODS[m];
        if (!(c->receivers & d->recv_bit)) continue;
        if (s->recv_hash[m] != recv) continue;
        recv_known = 1;
        if (s->method_hash[m] != method) continue;
        r->status = PTYPE_ST_OK;
        po_execute(d->opcode, a, check_limit, r);
        return;
    }
    r->status = recv_known ? PTYPE_ST_CANT_FIND_METHOD : PTYPE_ST_CANT_FIND_SERVICE;
}

void po_call(const po_snapshot* s, const char* service, uint64_t seq, const char* service_method,
             const int64_t args[3], uint64_t check_limit, ptype_reply* reply) {
    uint32_t method, recv;
    po_method_key(service_method, &method, &recv);
    po_call_client(s, po_find_name(s, service), seq, method, recv, args, check_limit, reply);
}

void po_dispatch(const po_snapshot* s, const ptype_req* reqs, size_t n, ptype_reply* replies,
                 uint64_t check_limit) {
    const po_client* last = NULL;
    for (size_t j = 0; j < n; j++) {
        const ptype_req* q = &reqs[j];
        const po_client* c = (last && last->key == q->svc_key) ? last : po_find_key(s, q->svc_key);
        if (c) last = c;
        int64_t a[3] = {q->a0, q->a1, q->a2};
        po_call_client(s, c, q->seq, q->method, q->recv, a, check_limit, &replies[j]);
    }
}

/* rpc.go:177-182 per record, in order:
 *   clients := c.getClients(); if len(clients) == 0 { return nil }
 *   index := atomic.AddUint64(&c.seq, uint64(1))                                  */
void po_assign_seq(const po_snapshot* s, ptype_req* reqs, size_t n, uint64_t* counters) {
    for (size_t j = 0; j < n; j++) {
        const po_client* c = po_find_key(s, reqs[j].svc_key);
        if (!c || c->k == 0) {
            reqs[j].seq = 0;
            continue;
        }
        reqs[j].seq = ++counters[c - s->clients];
    }
}
size_t po_snapshot_n_clients(const po_snapshot* s) { return s->n_clients; }
long po_snapshot_client_index(const po_snapshot* s, const char* service) {
    for (size_t j = 0; j < s->n_clients; j++)
        if (!strcmp(s->clients[j].name, service)) return (long)j;
    return -1;
}

typedef struct {
    const po_snapshot* s;
    const ptype_req* reqs;
    ptype_reply* replies;
    size_t n;
    uint64_t limit;
} po_job;
static void* po_worker(void* p) {
    po_job* j = (po_job*)p;
    po_dispatch(j->s, j->reqs, j->n, j->replies, j->limit);
    return NULL;
}

/* The timed CPU baseline: the batch split evenly over `threads` OS threads. */
void po_dispatch_mt(const po_snapshot* s, const ptype_req* reqs, size_t n, ptype_reply* replies,
                    uint64_t check_limit, int threads) {
    if (threads <= 1 || n < (size_t)threads * 64) {
        po_dispatch(s, reqs, n, replies, check_limit);
        return;
    }
    pthread_t* th = (pthread_t*)calloc((size_t)threads, sizeof *th);
    po_job* jobs = (po_job*)calloc((size_t)threads, sizeof *jobs);
    size_t per = (n + (size_t)threads - 1) / (size_t)threads;
    int started = 0;
    for (int t = 0; t < threads; t++) {
        size_t lo = (size_t)t * per;
        if (lo >= n) break;
        size_t hi = lo + per < n ? lo + per : n;
        jobs[t] = (po_job){s, reqs + lo, replies + lo, hi - lo, check_limit};
        pthread_create(&th[t], NULL, po_worker, &jobs[t]);
        started++;
    }
    for (int t = 0; t < started; t++) pthread_join(th[t], NULL);
    free(th);
    free(jobs);
}

/* Order-independent stream checksum (definition in include/ptype_b200.h;
 * no counterpart in the reference). */
static uint64_t po_fmix64(uint64_t x) {
    x ^= x >> 33;
    x *= 0xff51afd7ed558ccdull;
    x ^= x >> 33;
    x *= 0xc4ceb9fe1a85ec53ull;
    x ^= x >> 33;
    return x;
}
uint64_t po_reply_checksum(const ptype_reply* r, size_t n, uint64_t index_base) {
    uint64_t sum = 0;
    for (size_t j = 0; j < n; j++) {
        uint64_t idx = index_base + j;
        uint64_t w0 = (uint64_t)r[j].value;
        uint64_t w1 = (uint64_t)r[j].node | ((uint64_t)r[j].status << 32);
        sum += po_fmix64(w0 + 0x9E3779B97F4A7C15ull * (idx + 1)) ^
               po_fmix64(w1 ^ (idx * 0xC2B2AE3D27D4EB4Full + 0x165667B19E3779F9ull));
    }
    return sum;
}

/* ====================================================== encoding/gob values */
/* [Go stdlib encoding/gob; SURVEY.md appendix A]  unsigned: < 128 one byte; else
 * a byte holding the negated byte count, then minimal big-endian bytes. */
int po_gob_uint(uint64_t v, uint8_t* out) {
    if (v < 128) {
        out[0] = (uint8_t)v;
        return 1;
    }
    int nb = 0;
    for (uint64_t t = v; t; t >>= 8) nb++;
    out[0] = (uint8_t)(-nb);
    for (int i = 0; i < nb; i++) out[1 + i] = (uint8_t)(v >> (8 * (nb - 1 - i)));
    return nb + 1;
}
/* signed i: u = i<<1 if i >= 0 else (^i<<1)|1, then as unsigned. */
int po_gob_int(int64_t i, uint8_t* out) {
    uint64_t u = i < 0 ? ((~(uint64_t)i) << 1) | 1 : (uint64_t)i << 1;
    return po_gob_uint(u, out);
}
