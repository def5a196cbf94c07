/*
 * loopback.c — CPU restatement of the reference's REAL call path for BASELINE config 1
 * ("example/calculator single-node on CPU loopback"): cluster.(*Client).Call
 * (cluster/rpc.go:59-67) -> net/rpc client -> TCP 127.0.0.1 -> net/rpc server -> method.
 *
 * TEST INFRASTRUCTURE ONLY (a timed CPU baseline and a wire-format reference for the
 * gob value encodings); the product never links this file.
 *
 * What is restated, with its source:
 *   dial + handshake   rpc.DialHTTP (cluster/rpc.go:277): "CONNECT /_goRPC_ HTTP/1.0\n\n",
 *                      answered by rpc.HandleHTTP's handler (server.go:20) with
 *                      "HTTP/1.0 200 Connected to Go RPC\n\n"           [Go 1.13 net/rpc]
 *   per call, client   gob(Request{ServiceMethod, Seq}) + gob(args), one buffered flush;
 *                      then gob(Response{ServiceMethod, Seq, Error}) + gob(reply)
 *   per call, server   readRequestHeader: LastIndex(ServiceMethod, ".") -> serviceMap ->
 *                      method map; decode args; call; encode response      [net/rpc]
 *   node choice        roundRobinSelect over the dialled connections (cluster/rpc.go:176-183)
 *   method bodies      po_multiply / po_add / po_check (ptype_oracle.c)
 *   gob                uvarint length, type id, type definitions sent once per connection,
 *                      struct = (field delta, value)* 00 with zero fields omitted, top-level
 *                      non-struct values preceded by 00                    [encoding/gob]
 * The Go runtime itself (goroutine per request, reflection, mutexes, GC) is NOT restated,
 * so this is a lower bound on the Go path's cost, not a measurement of Go.
 * Type ids of user types are process-dependent in Go (SURVEY.md 8c iii); fixed here.
 */
#define _POSIX_C_SOURCE 200809L
#include <arpa/inet.h>
#include <errno.h>
#include <netinet/in.h>
#include <netinet/tcp.h>
#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/socket.h>
#include <time.h>
#include <unistd.h>

#include "ptype_oracle.h"

/* ------------------------------------------------------------------ buffers */
typedef struct {
    uint8_t* p;
    size_t n, cap;
} wbuf;

static void wb_put(wbuf* b, const void* src, size_t n) {
    if (b->n + n > b->cap) {
        b->cap = (b->n + n) * 2 + 64;
        b->p = (uint8_t*)realloc(b->p, b->cap);
    }
    memcpy(b->p + b->n, src, n);
    b->n += n;
}
static void wb_byte(wbuf* b, uint8_t v) { wb_put(b, &v, 1); }
static void wb_uint(wbuf* b, uint64_t v) {
    uint8_t t[9];
    wb_put(b, t, (size_t)po_gob_uint(v, t));
}
static void wb_int(wbuf* b, int64_t v) {
    uint8_t t[9];
    wb_put(b, t, (size_t)po_gob_int(v, t));
}
static void wb_string(wbuf* b, const char* s) {
    size_t l = strlen(s);
    wb_uint(b, l);
    wb_put(b, s, l);
}
/* one gob message: uvarint(len(body)) body */
static void wb_message(wbuf* out, const wbuf* body) {
    wb_uint(out, body->n);
    wb_put(out, body->p, body->n);
}

typedef struct {
    int fd;
    uint8_t buf[65536];
    size_t lo, hi;
    uint64_t total;
} rbuf; /* bufio.Reader */

static int rb_fill(rbuf* r) {
    if (r->lo == r->hi) r->lo = r->hi = 0;
    ssize_t k = read(r->fd, r->buf + r->hi, sizeof r->buf - r->hi);
    if (k <= 0) return -1;
    r->hi += (size_t)k;
    r->total += (uint64_t)k;
    return 0;
}
static int rb_byte(rbuf* r, uint8_t* v) {
    if (r->lo == r->hi && rb_fill(r)) return -1;
    *v = r->buf[r->lo++];
    return 0;
}
static int rb_read(rbuf* r, void* dst, size_t n) {
    uint8_t* d = (uint8_t*)dst;
    while (n) {
        if (r->lo == r->hi && rb_fill(r)) return -1;
        size_t k = r->hi - r->lo < n ? r->hi - r->lo : n;
        memcpy(d, r->buf + r->lo, k);
        r->lo += k;
        d += k;
        n -= k;
    }
    return 0;
}
static int rb_uint(rbuf* r, uint64_t* v) {
    uint8_t b;
    if (rb_byte(r, &b)) return -1;
    if (b < 128) {
        *v = b;
        return 0;
    }
    int nb = (int)(uint8_t)(-(int8_t)b);
    if (nb < 1 || nb > 8) return -1;
    uint64_t x = 0;
    for (int i = 0; i < nb; i++) {
        if (rb_byte(r, &b)) return -1;
        x = (x << 8) | b;
    }
    *v = x;
    return 0;
}
static int rb_int(rbuf* r, int64_t* v) {
    uint64_t u;
    if (rb_uint(r, &u)) return -1;
    *v = (u & 1) ? (int64_t) ~(u >> 1) : (int64_t)(u >> 1);
    return 0;
}

/* --------------------------------------------------------- gob type tables */
enum { T_INT = 2, T_UINT = 3, T_STRING = 6, T_REQUEST = 65, T_RESPONSE = 66, T_ARGS2 = 67, T_ARGS3 = 68, T_EMPTY = 69 };

/* wireType{StructT: &structType{CommonType{Name, Id}, Field: []fieldType{{Name, Id}...}}} */
static void gob_typedef_struct(wbuf* out, int id, const char* name, int nfields, const char* const* fnames,
                               const int* fids) {
    wbuf b = {0};
    wb_int(&b, -id);       /* negative id introduces a definition */
    wb_byte(&b, 3);        /* wireType: delta to StructT (field 2) */
    wb_byte(&b, 1);        /*   structType: delta to CommonType */
    wb_byte(&b, 1);        /*     CommonType.Name */
    wb_string(&b, name);
    wb_byte(&b, 1);        /*     CommonType.Id */
    wb_int(&b, id);
    wb_byte(&b, 0);        /*   end CommonType */
    if (nfields) {
        wb_byte(&b, 1);    /*   structType.Field */
        wb_uint(&b, (uint64_t)nfields);
        for (int i = 0; i < nfields; i++) {
            wb_byte(&b, 1);
            wb_string(&b, fnames[i]);
            wb_byte(&b, 1);
            wb_int(&b, fids[i]);
            wb_byte(&b, 0);
        }
    }
    wb_byte(&b, 0); /* end structType */
    wb_byte(&b, 0); /* end wireType */
    wb_message(out, &b);
    free(b.p);
}

/* Next VALUE message: skips type-definition messages; returns its type id and leaves the
 * reader at the value. */
static int gob_next_value(rbuf* r, int64_t* type_id) {
    for (;;) {
        uint64_t len;
        if (rb_uint(r, &len)) return -1;
        uint64_t before = r->total - (r->hi - r->lo);
        int64_t id;
        if (rb_int(r, &id)) return -1;
        if (id >= 0) {
            *type_id = id;
            return 0;
        }
        uint64_t used = (r->total - (r->hi - r->lo)) - before;
        uint8_t skip[256];
        uint64_t left = len - used;
        while (left) {
            size_t k = left < sizeof skip ? (size_t)left : sizeof skip;
            if (rb_read(r, skip, k)) return -1;
            left -= k;
        }
    }
}

static int gob_read_string(rbuf* r, char* out, size_t cap) {
    uint64_t l;
    if (rb_uint(r, &l) || l >= cap) return -1;
    if (rb_read(r, out, (size_t)l)) return -1;
    out[l] = 0;
    return 0;
}

/* ------------------------------------------------------------------ server */
typedef struct {
    int listen_fd;
    int port;
    pthread_t th;
    uint64_t check_limit;
    uint32_t receivers; /* bit0 Calculator, bit1 Prime: what this node rpc.Register-ed */
} lb_server;

static void send_all(int fd, const uint8_t* p, size_t n, uint64_t* counter) {
    while (n) {
        ssize_t k = write(fd, p, n);
        if (k <= 0) {
            if (errno == EINTR) continue;
            return;
        }
        p += k;
        n -= (size_t)k;
        if (counter) *counter += (uint64_t)k;
    }
}

/* What the server does with one decoded request [net/rpc server.readRequestHeader + service.call]:
 * split at the last '.', find the receiver among those this node registered, find the method,
 * run it.  Returns NULL and *reply on success, else the error text net/rpc would send. */
static const char* server_execute(const lb_server* srv, const char* sm, const int64_t* a, int64_t* reply,
                                  char* errbuf, size_t errcap) {
    const char* dot = strrchr(sm, '.');
    *reply = 0;
    if (!dot) {
        snprintf(errbuf, errcap, "rpc: service/method request ill-formed: %s", sm);
        return errbuf;
    }
    size_t rl = (size_t)(dot - sm);
    int is_calc = rl == 10 && !strncmp(sm, "Calculator", 10) && (srv->receivers & 1);
    int is_prime = rl == 5 && !strncmp(sm, "Prime", 5) && (srv->receivers & 2);
    if (!is_calc && !is_prime) {
        snprintf(errbuf, errcap, "rpc: can't find service %s", sm);
        return errbuf;
    }
    if (is_calc && !strcmp(dot + 1, "Multiply")) {
        *reply = po_multiply(a[0], a[1]);
    } else if (is_calc && !strcmp(dot + 1, "Add")) {
        *reply = po_add(a[0], a[1]);
    } else if (is_prime && !strcmp(dot + 1, "Check")) {
        if (po_check_width(a[0], a[1], a[2]) > srv->check_limit) {
            snprintf(errbuf, errcap, "ptype: check range beyond limit");
            return errbuf;
        }
        *reply = po_check(a[0], a[1], a[2]);
    } else {
        snprintf(errbuf, errcap, "rpc: can't find method %s", sm);
        return errbuf;
    }
    return NULL;
}

/* Per-connection encoder state of the server: which type definitions have gone out. */
typedef struct {
    int sent_types, sent_empty;
} lb_server_wire;

/* sendResponse [net/rpc]: gob(Response{ServiceMethod, Seq, Error}) then gob(reply), or on error
 * gob(invalidRequest = struct{}{}); type definitions precede their first use.  Appends to out. */
static void server_encode_response(lb_server_wire* st, const char* sm, uint64_t seq, const char* err,
                                   int64_t reply, wbuf* out) {
    if (!st->sent_types) {
        const char* fn[] = {"ServiceMethod", "Seq", "Error"};
        const int fi[] = {T_STRING, T_UINT, T_STRING};
        gob_typedef_struct(out, T_RESPONSE, "Response", 3, fn, fi);
        st->sent_types = 1;
    }
    wbuf b = {0};
    wb_int(&b, T_RESPONSE);
    wb_byte(&b, 1);
    wb_string(&b, sm);
    if (seq) {
        wb_byte(&b, 1);
        wb_uint(&b, seq);
    }
    if (err) {
        wb_byte(&b, (uint8_t)(seq ? 1 : 2));
        wb_string(&b, err);
    }
    wb_byte(&b, 0);
    wb_message(out, &b);
    b.n = 0;
    if (err) { /* body is invalidRequest = struct{}{} */
        if (!st->sent_empty) {
            gob_typedef_struct(out, T_EMPTY, "", 0, NULL, NULL);
            st->sent_empty = 1;
        }
        wb_int(&b, T_EMPTY);
        wb_byte(&b, 0);
    } else {
        wb_int(&b, T_INT);
        wb_byte(&b, 0); /* top-level non-struct marker */
        wb_int(&b, reply);
    }
    wb_message(out, &b);
    free(b.p);
}

/* ServeConn loop of one accepted connection [net/rpc server.ServeCodec]. */
static void serve_conn(int fd, const lb_server* srv) {
    rbuf* r = (rbuf*)calloc(1, sizeof *r);
    r->fd = fd;
    char line[256];
    size_t ll = 0;
    /* HTTP CONNECT handshake: read up to the blank line */
    for (;;) {
        uint8_t c;
        if (rb_byte(r, &c)) goto done;
        if (ll + 1 < sizeof line) line[ll++] = (char)c;
        if (ll >= 2 && line[ll - 1] == '\n' && line[ll - 2] == '\n') break;
    }
    line[ll] = 0;
    if (strncmp(line, "CONNECT /_goRPC_ ", 17) != 0) goto done;
    {
        const char* ok = "HTTP/1.0 200 Connected to Go RPC\n\n";
        send_all(fd, (const uint8_t*)ok, strlen(ok), NULL);
    }
    lb_server_wire wire = {0, 0};
    wbuf out = {0};
    for (;;) {
        /* ---- readRequestHeader */
        int64_t tid;
        if (gob_next_value(r, &tid)) break;
        char sm[256] = "";
        uint64_t seq = 0;
        int field = -1;
        for (;;) {
            uint64_t delta;
            if (rb_uint(r, &delta)) goto out_loop;
            if (delta == 0) break;
            field += (int)delta;
            if (field == 0) {
                if (gob_read_string(r, sm, sizeof sm)) goto out_loop;
            } else if (field == 1) {
                if (rb_uint(r, &seq)) goto out_loop;
            }
        }
        /* ---- args value: Args{A,B} or Args{Min,Max,Target}; zero fields are omitted */
        if (gob_next_value(r, &tid)) break;
        int64_t a[3] = {0, 0, 0};
        field = -1;
        for (;;) {
            uint64_t delta;
            if (rb_uint(r, &delta)) goto out_loop;
            if (delta == 0) break;
            field += (int)delta;
            int64_t v;
            if (rb_int(r, &v)) goto out_loop;
            if (field >= 0 && field < 3) a[field] = v;
        }
        /* ---- service / method lookup, the call, sendResponse */
        char errbuf[320];
        int64_t reply = 0;
        const char* err = server_execute(srv, sm, a, &reply, errbuf, sizeof errbuf);
        out.n = 0;
        server_encode_response(&wire, sm, seq, err, reply, &out);
        send_all(fd, out.p, out.n, NULL);
    }
out_loop:
    free(out.p);
done:
    free(r);
    close(fd);
}

typedef struct {
    int fd;
    const lb_server* srv;
} conn_arg;
static void* conn_main(void* p) {
    conn_arg* a = (conn_arg*)p;
    serve_conn(a->fd, a->srv);
    free(a);
    return NULL;
}

/* http.ListenAndServe + one goroutine per connection (server.go:38). */
static void* server_main(void* p) {
    lb_server* s = (lb_server*)p;
    for (;;) {
        int fd = accept(s->listen_fd, NULL, NULL);
        if (fd < 0) break;
        int one = 1;
        setsockopt(fd, IPPROTO_TCP, TCP_NODELAY, &one, sizeof one);
        conn_arg* a = (conn_arg*)malloc(sizeof *a);
        a->fd = fd;
        a->srv = s;
        pthread_t th;
        pthread_create(&th, NULL, conn_main, a);
        pthread_detach(th);
    }
    return NULL;
}

static int server_start(lb_server* s) {
    s->listen_fd = socket(AF_INET, SOCK_STREAM, 0);
    if (s->listen_fd < 0) return -1;
    int one = 1;
    setsockopt(s->listen_fd, SOL_SOCKET, SO_REUSEADDR, &one, sizeof one);
    struct sockaddr_in sa;
    memset(&sa, 0, sizeof sa);
    sa.sin_family = AF_INET;
    sa.sin_addr.s_addr = htonl(INADDR_LOOPBACK);
    sa.sin_port = 0; /* like the reference's tests: 127.0.0.1:0 (rpc_test.go:97) */
    if (bind(s->listen_fd, (struct sockaddr*)&sa, sizeof sa) || listen(s->listen_fd, 64)) return -1;
    socklen_t sl = sizeof sa;
    getsockname(s->listen_fd, (struct sockaddr*)&sa, &sl);
    s->port = ntohs(sa.sin_port);
    return pthread_create(&s->th, NULL, server_main, s);
}

/* ------------------------------------------------------------------ client */
/* Per-connection encoder state of the client. */
typedef struct {
    uint64_t rpc_seq; /* net/rpc's own Seq, starts at 0 per connection */
    int sent_request_type, sent_args2, sent_args3;
} lb_client_wire;

typedef struct {
    int fd;
    rbuf* r;
    lb_client_wire wire;
    uint64_t bytes_out;
} lb_conn;

static int conn_dial(lb_conn* c, int port) { /* rpc.DialHTTP (rpc.go:277) */
    memset(c, 0, sizeof *c);
    c->fd = socket(AF_INET, SOCK_STREAM, 0);
    if (c->fd < 0) return -1;
    int one = 1;
    setsockopt(c->fd, IPPROTO_TCP, TCP_NODELAY, &one, sizeof one);
    struct sockaddr_in sa;
    memset(&sa, 0, sizeof sa);
    sa.sin_family = AF_INET;
    sa.sin_addr.s_addr = htonl(INADDR_LOOPBACK);
    sa.sin_port = htons((uint16_t)port);
    if (connect(c->fd, (struct sockaddr*)&sa, sizeof sa)) return -1;
    const char* hello = "CONNECT /_goRPC_ HTTP/1.0\n\n";
    send_all(c->fd, (const uint8_t*)hello, strlen(hello), &c->bytes_out);
    c->r = (rbuf*)calloc(1, sizeof(rbuf));
    c->r->fd = c->fd;
    char line[128];
    size_t ll = 0;
    for (;;) {
        uint8_t ch;
        if (rb_byte(c->r, &ch)) return -1;
        if (ll + 1 < sizeof line) line[ll++] = (char)ch;
        if (ll >= 2 && line[ll - 1] == '\n' && line[ll - 2] == '\n') break;
    }
    line[ll] = 0;
    return strncmp(line, "HTTP/1.0 200 Connected to Go RPC", 32) ? -1 : 0;
}

/* (*rpc.Client).Call: one request in flight, wait for its response. */
/* What (*rpc.Client).send writes for one call [net/rpc + encoding/gob]: gob(Request{ServiceMethod,
 * Seq}) then gob(args); type definitions precede their first use on the connection.  Appends. */
static void client_encode_call(lb_client_wire* st, const char* sm, const int64_t* a, int nargs, wbuf* out) {
    if (!st->sent_request_type) {
        const char* fn[] = {"ServiceMethod", "Seq"};
        const int fi[] = {T_STRING, T_UINT};
        gob_typedef_struct(out, T_REQUEST, "Request", 2, fn, fi);
        st->sent_request_type = 1;
    }
    wbuf b = {0};
    wb_int(&b, T_REQUEST);
    wb_byte(&b, 1);
    wb_string(&b, sm);
    if (st->rpc_seq) {
        wb_byte(&b, 1);
        wb_uint(&b, st->rpc_seq);
    }
    wb_byte(&b, 0);
    wb_message(out, &b);
    st->rpc_seq++;
    int tid = nargs == 3 ? T_ARGS3 : T_ARGS2;
    if (nargs == 3 && !st->sent_args3) {
        const char* fn[] = {"Min", "Max", "Target"};
        const int fi[] = {T_INT, T_INT, T_INT};
        gob_typedef_struct(out, T_ARGS3, "Args", 3, fn, fi);
        st->sent_args3 = 1;
    } else if (nargs != 3 && !st->sent_args2) {
        const char* fn[] = {"A", "B"};
        const int fi[] = {T_INT, T_INT};
        gob_typedef_struct(out, T_ARGS2, "Args", 2, fn, fi);
        st->sent_args2 = 1;
    }
    b.n = 0;
    wb_int(&b, tid);
    int last = -1;
    for (int f = 0; f < nargs; f++) {
        if (a[f] == 0) continue; /* zero-valued fields are not sent */
        wb_byte(&b, (uint8_t)(f - last));
        wb_int(&b, a[f]);
        last = f;
    }
    wb_byte(&b, 0);
    wb_message(out, &b);
    free(b.p);
}

static int conn_call(lb_conn* c, const char* sm, const int64_t* a, int nargs, ptype_reply* rep, wbuf* out) {
    out->n = 0;
    client_encode_call(&c->wire, sm, a, nargs, out);
    send_all(c->fd, out->p, out->n, &c->bytes_out);

    /* ---- input(): Response then body */
    int64_t t;
    if (gob_next_value(c->r, &t)) return -1;
    char rsm[256] = "", err[320] = "";
    uint64_t seq = 0;
    int field = -1;
    for (;;) {
        uint64_t delta;
        if (rb_uint(c->r, &delta)) return -1;
        if (delta == 0) break;
        field += (int)delta;
        if (field == 0) {
            if (gob_read_string(c->r, rsm, sizeof rsm)) return -1;
        } else if (field == 1) {
            if (rb_uint(c->r, &seq)) return -1;
        } else if (field == 2) {
            if (gob_read_string(c->r, err, sizeof err)) return -1;
        }
    }
    if (gob_next_value(c->r, &t)) return -1;
    rep->value = 0;
    if (err[0]) {
        uint8_t z;
        if (rb_byte(c->r, &z)) return -1; /* struct{} value: just the terminator */
        if (!strncmp(err, "rpc: service/method request ill-formed", 38))
            rep->status = PTYPE_ST_ILL_FORMED;
        else if (!strncmp(err, "rpc: can't find service", 23))
            rep->status = PTYPE_ST_CANT_FIND_SERVICE;
        else if (!strncmp(err, "rpc: can't find method", 22))
            rep->status = PTYPE_ST_CANT_FIND_METHOD;
        else
            rep->status = PTYPE_ST_RANGE_LIMIT;
    } else {
        uint8_t z;
        int64_t v;
        if (rb_byte(c->r, &z) || rb_int(c->r, &v)) return -1;
        rep->value = v;
        rep->status = PTYPE_ST_OK;
    }
    return 0;
}

/* ---------------------------------------------------------------- the run */
static const char* method_string(uint32_t method_hash, int* nargs) {
    static const char* names[] = {"Calculator.Multiply", "Calculator.Add", "Prime.Check"};
    static uint32_t h[3];
    static int init = 0;
    if (!init) {
        for (int i = 0; i < 3; i++) h[i] = po_fnv1a32(names[i], strlen(names[i]));
        init = 1;
    }
    for (int i = 0; i < 3; i++)
        if (h[i] == method_hash) {
            *nargs = i == 2 ? 3 : 2;
            return names[i];
        }
    *nargs = 2;
    return NULL;
}

/* N sequential synchronous Client.Call()s of ONE Client whose balancer holds `k`
 * connections (one server per selected node, every server registers `receivers`).
 * Request j goes to connection int(seq_j) % k (rpc.go:181-182); replies[j].node is
 * that connection index.  Requests whose method hash is none of the three known strings
 * are sent as "Unknown.Method".  Returns 0 on success. */
int po_loopback_run(const ptype_req* reqs, size_t n, ptype_reply* replies, int k, uint32_t receivers,
                    uint64_t check_limit, po_loopback_stats* stats) {
    if (k < 1 || k > 64) return -1;
    lb_server* srv = (lb_server*)calloc((size_t)k, sizeof *srv);
    lb_conn* conns = (lb_conn*)calloc((size_t)k, sizeof *conns);
    int rc = 0;
    for (int i = 0; i < k; i++) {
        srv[i].check_limit = check_limit;
        srv[i].receivers = receivers;
        if (server_start(&srv[i])) rc = -1;
    }
    for (int i = 0; i < k && !rc; i++)
        if (conn_dial(&conns[i], srv[i].port)) rc = -1; /* connectToNodes (rpc.go:272-285) */
    wbuf out = {0};
    struct timespec t0, t1;
    clock_gettime(CLOCK_MONOTONIC, &t0);
    for (size_t j = 0; j < n && !rc; j++) {
        int64_t idx = po_round_robin(reqs[j].seq, (size_t)k);
        if (idx < 0) {
            replies[j].value = 0;
            replies[j].node = PTYPE_NO_NODE;
            replies[j].status = PTYPE_ST_BAD_INDEX;
            continue;
        }
        int nargs;
        const char* sm = method_string(reqs[j].method, &nargs);
        if (!sm) sm = reqs[j].recv == PTYPE_RECV_ILL_FORMED ? "nodot" : "Unknown.Method";
        int64_t a[3] = {reqs[j].a0, reqs[j].a1, reqs[j].a2};
        replies[j].node = (uint32_t)idx;
        if (conn_call(&conns[idx], sm, a, nargs, &replies[j], &out)) rc = -2;
    }
    clock_gettime(CLOCK_MONOTONIC, &t1);
    if (stats) {
        stats->seconds = (double)(t1.tv_sec - t0.tv_sec) + 1e-9 * (double)(t1.tv_nsec - t0.tv_nsec);
        stats->calls = n;
        stats->bytes_client_to_server = 0;
        stats->bytes_server_to_client = 0;
        for (int i = 0; i < k; i++) {
            stats->bytes_client_to_server += conns[i].bytes_out;
            if (conns[i].r) stats->bytes_server_to_client += conns[i].r->total;
        }
    }
    for (int i = 0; i < k; i++) {
        if (conns[i].fd > 0) close(conns[i].fd);
        free(conns[i].r);
        if (srv[i].listen_fd > 0) {
            shutdown(srv[i].listen_fd, SHUT_RDWR);
            close(srv[i].listen_fd);
            pthread_join(srv[i].th, NULL);
        }
    }
    free(out.p);
    free(srv);
    free(conns);
    return rc;
}

/* The bytes ONE fresh connection carries for the calls reqs[0..n), after the HTTP CONNECT
 * handshake: what the client writes (c2s) and what the server answers (s2c), produced by the very
 * encoders po_loopback_run sends over TCP.  Either buffer may be NULL.  off[j] (j = 0..n) is
 * where call j's first byte lies (type definitions sent with a call count as part of it).
 * Method strings are chosen from the records like po_loopback_run does.  Returns 0, or -1 when a
 * buffer is too small (sizes needed are still reported through off[n]). */
int po_wire_streams(const ptype_req* reqs, size_t n, uint32_t receivers, uint64_t check_limit, uint8_t* c2s,
                    size_t c2s_cap, uint64_t* c2s_off, uint8_t* s2c, size_t s2c_cap, uint64_t* s2c_off) {
    lb_client_wire cw;
    lb_server_wire sw = {0, 0};
    lb_server srv;
    memset(&cw, 0, sizeof cw);
    memset(&srv, 0, sizeof srv);
    srv.receivers = receivers;
    srv.check_limit = check_limit;
    wbuf a = {0}, b = {0};
    for (size_t j = 0; j < n; j++) {
        int nargs;
        const char* sm = method_string(reqs[j].method, &nargs);
        if (!sm) sm = reqs[j].recv == PTYPE_RECV_ILL_FORMED ? "nodot" : "Unknown.Method";
        const int64_t args[3] = {reqs[j].a0, reqs[j].a1, nargs == 3 ? reqs[j].a2 : 0};
        if (c2s_off) c2s_off[j] = a.n;
        if (s2c_off) s2c_off[j] = b.n;
        const uint64_t seq = cw.rpc_seq;
        client_encode_call(&cw, sm, args, nargs, &a);
        char errbuf[320];
        int64_t reply = 0;
        const char* err = server_execute(&srv, sm, args, &reply, errbuf, sizeof errbuf);
        server_encode_response(&sw, sm, seq, err, reply, &b);
    }
    if (c2s_off) c2s_off[n] = a.n;
    if (s2c_off) s2c_off[n] = b.n;
    int rc = 0;
    if (c2s) {
        if (a.n <= c2s_cap) memcpy(c2s, a.p, a.n); else rc = -1;
    }
    if (s2c) {
        if (b.n <= s2c_cap) memcpy(s2c, b.p, b.n); else rc = -1;
    }
    free(a.p);
    free(b.p);
    return rc;
}

/* The wire bytes of one call on a fresh connection, for tests: request side into
 * `req_out`, and what a server would answer into `resp_out`.  Returns lengths. */
int po_gob_reply_int_message(int64_t v, uint8_t* out) {
    wbuf b = {0}, m = {0};
    wb_int(&b, T_INT);
    wb_byte(&b, 0);
    wb_int(&b, v);
    wb_message(&m, &b);
    int n = (int)m.n;
    memcpy(out, m.p, m.n);
    free(b.p);
    free(m.p);
    return n;
}

int po_gob_args_body(const int64_t* a, int nargs, uint8_t* out) {
    wbuf b = {0};
    int last = -1;
    for (int f = 0; f < nargs; f++) {
        if (a[f] == 0) continue;
        wb_byte(&b, (uint8_t)(f - last));
        wb_int(&b, a[f]);
        last = f;
    }
    wb_byte(&b, 0);
    int n = (int)b.n;
    memcpy(out, b.p, b.n);
    free(b.p);
    return n;
}

int po_gob_request_typedef(uint8_t* out) {
    wbuf m = {0};
    const char* fn[] = {"ServiceMethod", "Seq"};
    const int fi[] = {T_STRING, T_UINT};
    gob_typedef_struct(&m, T_REQUEST, "Request", 2, fn, fi);
    int n = (int)m.n;
    memcpy(out, m.p, m.n);
    free(m.p);
    return n;
}
