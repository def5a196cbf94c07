// wire.cu — the byte streams of Go's net/rpc either side of the dispatch path, on the device
// (call sites of the wire in the reference: cluster/rpc.go:65,88,277 and
// example/calculator/server/server.go:16-20).
//
// A net/rpc connection carries, after the HTTP CONNECT handshake, a stream of encoding/gob
// messages: `uvarint length, signed type id, body`.  A NEGATIVE type id introduces a type
// definition (sent once per type and connection, before its first use); the client sends per call
// one rpc.Request{ServiceMethod string; Seq uint64} value and one args value, the server answers
// with one rpc.Response{ServiceMethod; Seq; Error string} value and one reply value (an empty
// struct when Error is set).  Structs are (field delta, value)* 00 with zero fields omitted;
// unsigned ints are one byte below 128, else a negated byte count and big-endian bytes; signed
// ints are zig-zag folded into unsigned ones; a top-level non-struct value is preceded by 00.
// [Go 1.13 encoding/gob, net/rpc; SURVEY.md appendix A.]
//
// Here: the client-to-server streams of many connections are framed and decoded straight into
// ptype_req records (wire_decode), and reply records are encoded into the server-to-client
// streams (wire_encode).  One warp owns one connection's stream, because a stream is sequential
// by construction (a message starts where the previous one ends, and the type definitions sent
// so far are per-connection state): lane 0 walks the length prefixes of the next 32 calls, then
// the 32 lanes decode or encode one call each.  Parallelism comes from the number of connections,
// which is what a server has.  Byte work, HBM-bound at best, latency-bound on few streams.
#include <cuda_runtime.h>
#include <stdint.h>

#include <string>

#include "../../include/ptype_b200.h"
#include "device_image.h"

namespace ptype {
namespace {

constexpr int kWireThreads = 128;  // four streams per CTA
constexpr uint32_t kFull = 0xFFFFFFFFu;

// ---------------------------------------------------------------- gob readers
// A stream is read through a window: the warp copies the next kWindow bytes of its stream into
// shared memory with coalesced 16-byte loads, and the byte-by-byte readers below (lane 0 walking
// length prefixes, then every lane parsing its own call) hit shared memory instead of chasing
// dependent global loads.  Bytes outside the window (a chunk of calls longer than the window) are
// still read from global memory.
constexpr uint32_t kWindow = 2048;
struct Cursor {
    const uint8_t* p;
    uint64_t pos, end;
    bool ok;
    const uint8_t* win;  // shared-memory copy of p[wbase .. wbase + wlen)
    uint64_t wbase;
    uint32_t wlen;
};
__device__ __forceinline__ uint32_t rd_byte(Cursor& c) {
    if (c.pos >= c.end) {
        c.ok = false;
        return 0;
    }
    const uint64_t off = c.pos - c.wbase;  // wraps to a huge value below the window
    const uint64_t at = c.pos++;
    return off < c.wlen ? c.win[off] : __ldg(c.p + at);
}
__device__ __forceinline__ uint64_t rd_uint(Cursor& c) {
    const uint32_t b = rd_byte(c);
    if (b < 128) return b;
    const uint32_t nb = 256u - b;
    if (nb < 1 || nb > 8) {
        c.ok = false;
        return 0;
    }
    uint64_t x = 0;
    for (uint32_t k = 0; k < nb; ++k) x = (x << 8) | rd_byte(c);
    return x;
}
__device__ __forceinline__ int64_t unzigzag(uint64_t u) {
    return (u & 1) ? static_cast<int64_t>(~(u >> 1)) : static_cast<int64_t>(u >> 1);
}

// One call as lane 0 framed it: where the bodies of its two value messages lie.
struct Framed {
    uint64_t hdr_pos, hdr_end, arg_pos, arg_end;
};

// Lane 0: walk messages from `c.pos` until one more call (a Request value and an args value, type
// definitions skipped) is complete.  False at the end of the stream or on a framing error.
// The walk is one lane running serial code, so its instruction count is what the whole decoder
// costs (the first version spent ~250 warp instructions per call here and saturated the SMs'
// issue slots with 1/32 of the lanes): the usual message - a one-byte length, a type id of at most
// nine bytes, everything inside the window - takes three shared-memory bytes and a few compares,
// because framing needs only the length and the SIGN of the type id (the low bit of its last byte).
__device__ bool next_call(Cursor& c, Framed& f) {
    int have = 0;
    while (c.ok && c.pos < c.end) {
        uint64_t body, len, value_pos;
        bool definition;
        const uint64_t off = c.pos - c.wbase;
        const uint32_t b0 = off + 11 <= c.wlen ? c.win[off] : 255u;
        if (b0 < 128) {  // fast path: 1-byte length, the id's bytes in the window
            len = b0;
            body = c.pos + 1;
            const uint32_t b1 = c.win[off + 1];
            uint32_t id_bytes = 1, last = b1;
            if (b1 >= 128) {
                const uint32_t nb = 256u - b1;
                if (nb > 8) {
                    c.ok = false;
                    return false;
                }
                id_bytes = 1 + nb;
                last = c.win[off + 1 + nb];
            }
            definition = (last & 1u) != 0;  // zig-zag: odd = negative = a type definition
            value_pos = body + id_bytes;
            if (len > c.end - body || id_bytes > len) {
                c.ok = false;
                return false;
            }
        } else {
            len = rd_uint(c);
            body = c.pos;
            if (!c.ok || len > c.end - body) {
                c.ok = false;
                return false;
            }
            Cursor m = c;
            m.pos = body;
            m.end = body + len;
            definition = unzigzag(rd_uint(m)) < 0;
            value_pos = m.pos;
            if (!m.ok) {
                c.ok = false;
                return false;
            }
        }
        c.pos = body + len;
        if (definition) continue;  // nothing to decode: the values are positional
        if (have == 0) {
            f.hdr_pos = value_pos;
            f.hdr_end = body + len;
            have = 1;
        } else {
            f.arg_pos = value_pos;
            f.arg_end = body + len;
            return true;
        }
    }
    if (have) c.ok = false;  // a header without its args
    return false;
}

// ------------------------------------------------------------------- decode
// kCount: only count the calls of every stream (calls[s]); else decode them into
// reqs/calls at call_off[s].
template <bool kCount>
__global__ void __launch_bounds__(kWireThreads)
    wire_decode_kernel(const uint8_t* __restrict__ bytes, const unsigned long long* __restrict__ stream_off,
                       const size_t n_streams, const ptype_wire_stream* __restrict__ streams,
                       unsigned long long* __restrict__ call_off, ptype_req* __restrict__ reqs,
                       ptype_wire_call* __restrict__ calls) {
    __shared__ Framed s_framed[kWireThreads / 32][32];
    __shared__ __align__(16) uint8_t s_window[kWireThreads / 32][kWindow];
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const size_t s = static_cast<size_t>(blockIdx.x) * (kWireThreads / 32) + warp;
    if (s >= n_streams) return;
    Cursor c = {bytes, stream_off[s], stream_off[s + 1], true, s_window[warp], 0, 0};
    unsigned long long next_pos = c.pos;  // where the next chunk of calls starts (lane 0 knows, all lanes need it)
    unsigned long long done = 0;
    const unsigned long long out0 = kCount ? 0ull : call_off[s];
    ptype_wire_stream st = {};
    if (!kCount) st = streams[s];
    for (;;) {
        // refill the window from the first byte not yet framed (16-byte chunks that lie wholly inside
        // the stream; `bytes` comes from cudaMalloc, so a 16-aligned offset is a 16-aligned address)
        {
            const uint64_t wbase = next_pos & ~15ull;
            const uint64_t avail = c.end > wbase ? c.end - wbase : 0;
            const uint32_t wlen = avail >= kWindow ? kWindow : static_cast<uint32_t>(avail & ~15ull);
            __syncwarp();  // the previous chunk's readers are done with the old window
            for (uint32_t o = lane * 16; o < wlen; o += 32 * 16)
                *reinterpret_cast<uint4*>(s_window[warp] + o) = __ldg(reinterpret_cast<const uint4*>(bytes + wbase + o));
            __syncwarp();
            c.wbase = wbase;
            c.wlen = wlen;
        }
        uint32_t n = 0;
        if (lane == 0) {
            // The usual message, walked in 32-bit offsets from the window's base: a one-byte length,
            // a type id of one byte or of FF + one byte (ids 64..127: every user type of a
            // connection), all inside the window.  Three independent shared-memory bytes and ~25
            // instructions per message; anything else - and the stream's end - goes to next_call,
            // from the start of the call in progress.  This loop is the decoder's cost: it runs on
            // one lane of 32.
            const uint64_t span = c.end - c.wbase;
            const uint32_t lim = span > 0xFFFFFFFFull ? 0xFFFFFFFFu : static_cast<uint32_t>(span);
            const uint8_t* const win = s_window[warp];
            uint32_t off = static_cast<uint32_t>(next_pos - c.wbase), call_off = off, hdr_v = 0, hdr_e = 0;
            bool have = false;
            while (n < 32 && off + 12 <= c.wlen) {
                const uint32_t b0 = win[off], b1 = win[off + 1], b2 = win[off + 2];
                const bool wide = b1 >= 128;
                const uint32_t id_bytes = wide ? 2u : 1u, last = wide ? b2 : b1;
                const uint32_t next = off + 1 + b0;
                if (b0 >= 128 || (wide && b1 != 0xFFu) || id_bytes > b0 || next > lim) break;
                if (!(last & 1u)) {  // a value, not a type definition
                    const uint32_t value = off + 1 + id_bytes;
                    if (!have) {
                        hdr_v = value;
                        hdr_e = next;
                        have = true;
                    } else {
                        if (!kCount)
                            s_framed[warp][n] = Framed{c.wbase + hdr_v, c.wbase + hdr_e, c.wbase + value, c.wbase + next};
                        ++n;
                        have = false;
                        call_off = next;
                    }
                }
                off = next;
            }
            c.pos = c.wbase + call_off;
            Framed f;
            while (n < 32 && next_call(c, f)) s_framed[warp][n++] = f;
            next_pos = c.pos;
        }
        n = __shfl_sync(kFull, n, 0);
        next_pos = __shfl_sync(kFull, next_pos, 0);
        __syncwarp();
        if (n == 0) break;
        if (!kCount && lane < n) {
            const Framed f = s_framed[warp][lane];
            // ---- rpc.Request{ServiceMethod, Seq}
            Cursor h = c;
            h.pos = f.hdr_pos;
            h.end = f.hdr_end;
            h.ok = true;
            uint32_t hash = 0x811c9dc5u, hash_at_dot = 0;
            bool dot = false;
            unsigned long long rpc_seq = 0, sm_off = 0;
            uint32_t sm_len = 0;
            int field = -1;
            for (;;) {
                const uint64_t delta = rd_uint(h);
                if (!h.ok || delta == 0) break;
                field += static_cast<int>(delta);
                if (field == 0) {
                    const uint64_t l = rd_uint(h);
                    if (!h.ok || l > h.end - h.pos) {
                        h.ok = false;
                        break;
                    }
                    sm_off = h.pos;
                    sm_len = static_cast<uint32_t>(l);
                    for (uint64_t k = 0; k < l; ++k) {  // FNV-1a-32, split at the LAST '.' like net/rpc's server
                        const uint32_t b = rd_byte(h);
                        if (b == '.') {
                            dot = true;
                            hash_at_dot = hash;
                        }
                        hash = (hash ^ b) * 0x01000193u;
                    }
                } else if (field == 1) {
                    rpc_seq = rd_uint(h);
                } else {
                    h.ok = false;
                    break;
                }
            }
            // ---- the args value: int fields by position (calculator.Args{A,B} / prime.Args{Min,Max,Target})
            Cursor a = c;
            a.pos = f.arg_pos;
            a.end = f.arg_end;
            a.ok = true;
            int64_t v[3] = {0, 0, 0};
            field = -1;
            for (;;) {
                const uint64_t delta = rd_uint(a);
                if (!a.ok || delta == 0) break;
                field += static_cast<int>(delta);
                const int64_t x = unzigzag(rd_uint(a));
                if (field >= 0 && field < 3) v[field] = x;
            }
            const unsigned long long j = done + lane;
            ptype_req q;
            q.svc_key = st.svc_key;
            q.seq = st.seq0 + j * st.seq_step;
            q.method = hash;
            q.recv = dot ? hash_at_dot : PTYPE_RECV_ILL_FORMED;
            q.a0 = v[0];
            q.a1 = v[1];
            q.a2 = v[2];
            reqs[out0 + j] = q;
            ptype_wire_call wc;
            wc.rpc_seq = rpc_seq;
            wc.sm_off = sm_off;
            wc.sm_len = sm_len;
            wc.flags = (h.ok && a.ok) ? 0u : PTYPE_WIRE_MALFORMED;
            calls[out0 + j] = wc;
        }
        done += n;
        __syncwarp();
    }
    if (kCount && lane == 0) call_off[s] = done;
}

// ------------------------------------------------------------------- encode
__device__ __forceinline__ uint32_t uint_len(uint64_t u) {
    if (u < 128) return 1;
    return 1 + ((64 - __clzll(static_cast<long long>(u))) + 7) / 8;
}
__device__ __forceinline__ uint64_t zigzag(int64_t i) {
    return i < 0 ? ((~static_cast<uint64_t>(i)) << 1) | 1ull : static_cast<uint64_t>(i) << 1;
}
__device__ __forceinline__ void put_uint(uint8_t*& p, uint64_t u) {
    if (u < 128) {
        *p++ = static_cast<uint8_t>(u);
        return;
    }
    const uint32_t nb = uint_len(u) - 1;
    *p++ = static_cast<uint8_t>(0u - nb);
    for (uint32_t k = 0; k < nb; ++k) *p++ = static_cast<uint8_t>(u >> (8 * (nb - 1 - k)));
}

// The Error text net/rpc's server puts into the Response for a status (the method string follows
// the first three).  [Go 1.13 net/rpc server.go: readRequestHeader]
__constant__ char kErrIllFormed[] = "rpc: service/method request ill-formed: ";
__constant__ char kErrService[] = "rpc: can't find service ";
__constant__ char kErrMethod[] = "rpc: can't find method ";
__constant__ char kErrRange[] = "ptype: check range beyond limit";
__constant__ char kErrOther[] = "ptype: call failed";

struct ErrText {
    const char* text;
    uint32_t len;
    bool with_method;
};
__device__ __forceinline__ ErrText err_text(uint32_t status) {
    switch (status) {
        case PTYPE_ST_ILL_FORMED: return {kErrIllFormed, sizeof kErrIllFormed - 1, true};
        case PTYPE_ST_CANT_FIND_SERVICE: return {kErrService, sizeof kErrService - 1, true};
        case PTYPE_ST_CANT_FIND_METHOD: return {kErrMethod, sizeof kErrMethod - 1, true};
        case PTYPE_ST_RANGE_LIMIT: return {kErrRange, sizeof kErrRange - 1, false};
        default: return {kErrOther, sizeof kErrOther - 1, false};
    }
}

struct WireTypes {
    const uint8_t* defs;     // the Response type definition message, then the empty struct's
    uint32_t response_def_len, empty_def_len;
    uint32_t response_id, empty_id;
};

// kSize: only the byte count of every stream (out_off[s]); else the bytes at out_off[s].
constexpr uint32_t kEncodeStage = 3072;  // bytes of one 32-call chunk assembled in shared memory
template <bool kSize>
__global__ void __launch_bounds__(kWireThreads)
    wire_encode_kernel(const uint8_t* __restrict__ req_bytes, const unsigned long long* __restrict__ call_off,
                       const size_t n_streams, const ptype_wire_call* __restrict__ calls,
                       const ptype_reply* __restrict__ replies, const WireTypes types, uint8_t* __restrict__ state,
                       unsigned long long* __restrict__ out_off, uint8_t* __restrict__ out) {
    __shared__ __align__(16) uint8_t s_stage[kSize ? 1 : kWireThreads / 32][kSize ? 16 : kEncodeStage];
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const size_t s = static_cast<size_t>(blockIdx.x) * (kWireThreads / 32) + warp;
    if (s >= n_streams) return;
    const unsigned long long first = call_off[s], last = call_off[s + 1];
    uint32_t st = state[s];  // bit 0: Response definition sent, bit 1: empty struct's definition sent
    unsigned long long at = kSize ? 0ull : out_off[s];
    for (unsigned long long base = first; base < last; base += 32) {
        const unsigned long long j = base + lane;
        const bool live = j < last;
        ptype_wire_call wc = {};
        ptype_reply rep = {};
        if (live) {
            wc = calls[j];
            rep = replies[j];
        }
        const bool failed = live && rep.status != PTYPE_ST_OK;
        const uint32_t fail_mask = __ballot_sync(kFull, failed);
        const bool def_response = live && lane == 0 && !(st & 1u);
        const bool def_empty = failed && !(st & 2u) && (fail_mask & ((1u << lane) - 1u)) == 0;
        const ErrText et = err_text(rep.status);
        const uint32_t err_len = failed ? et.len + (et.with_method ? wc.sm_len : 0u) : 0u;
        const uint64_t zz = zigzag(rep.value);
        // ---- sizes
        const uint32_t hdr_body = uint_len(zigzag(types.response_id)) + 1 + uint_len(wc.sm_len) + wc.sm_len +
                                  (wc.rpc_seq ? 1 + uint_len(wc.rpc_seq) : 0) +
                                  (failed ? 1 + uint_len(err_len) + err_len : 0) + 1;
        const uint32_t val_body = failed ? uint_len(zigzag(types.empty_id)) + 1 : 2 + uint_len(zz);
        uint32_t total = 0;
        if (live)
            total = (def_response ? types.response_def_len : 0) + uint_len(hdr_body) + hdr_body +
                    (def_empty ? types.empty_def_len : 0) + uint_len(val_body) + val_body;
        uint32_t incl = total;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t = __shfl_up_sync(kFull, incl, o);
            if (lane >= static_cast<uint32_t>(o)) incl += t;
        }
        const uint32_t chunk_total = __shfl_sync(kFull, incl, 31);
        if (!kSize) {
            // The chunk's bytes are contiguous in the stream (~36 per call).  They are assembled in
            // shared memory at the output's own 16-byte phase and leave as 16-byte stores, one run
            // per warp; written straight to global memory each lane's single-byte stores scattered
            // over 32 sectors per instruction.  A chunk too long for the stage (long method names
            // or error texts) still goes the direct way.
            uint8_t* const dst = out + at;
            const uint32_t phase = static_cast<uint32_t>(reinterpret_cast<uintptr_t>(dst) & 15u);
            const bool staged = chunk_total + phase <= kEncodeStage;  // warp-uniform
            if (live) {
                uint8_t* p = (staged ? s_stage[warp] + phase : dst) + (incl - total);
                if (def_response)
                    for (uint32_t k = 0; k < types.response_def_len; ++k) *p++ = __ldg(types.defs + k);
                // Response{ServiceMethod, Seq, Error}
                put_uint(p, hdr_body);
                put_uint(p, zigzag(types.response_id));
                *p++ = 1;
                put_uint(p, wc.sm_len);
                for (uint32_t k = 0; k < wc.sm_len; ++k) *p++ = __ldg(req_bytes + wc.sm_off + k);
                if (wc.rpc_seq) {
                    *p++ = 1;
                    put_uint(p, wc.rpc_seq);
                }
                if (failed) {
                    *p++ = wc.rpc_seq ? 1 : 2;
                    put_uint(p, err_len);
                    for (uint32_t k = 0; k < et.len; ++k) *p++ = static_cast<uint8_t>(et.text[k]);
                    if (et.with_method)
                        for (uint32_t k = 0; k < wc.sm_len; ++k) *p++ = __ldg(req_bytes + wc.sm_off + k);
                }
                *p++ = 0;
                // the reply value, or net/rpc's invalidRequest (struct{}{}) behind an error
                if (def_empty)
                    for (uint32_t k = 0; k < types.empty_def_len; ++k) *p++ = __ldg(types.defs + types.response_def_len + k);
                put_uint(p, val_body);
                if (failed) {
                    put_uint(p, zigzag(types.empty_id));
                    *p++ = 0;
                } else {
                    *p++ = 0x04;  // type id 2 (int)
                    *p++ = 0x00;  // top-level non-struct marker
                    put_uint(p, zz);
                }
            }
            if (staged) {
                __syncwarp();
                const uint8_t* const src = s_stage[warp] + phase;
                const uint32_t head = chunk_total < ((16u - phase) & 15u) ? chunk_total : ((16u - phase) & 15u);
                const uint32_t body = (chunk_total - head) & ~15u;
                if (lane < head) dst[lane] = src[lane];
                for (uint32_t o = lane * 16; o < body; o += 32 * 16)
                    *reinterpret_cast<uint4*>(dst + head + o) = *reinterpret_cast<const uint4*>(src + head + o);
                const uint32_t tail_at = head + body;
                if (tail_at + lane < chunk_total) dst[tail_at + lane] = src[tail_at + lane];
                __syncwarp();  // the stage is free for the next chunk
            }
        }
        at += chunk_total;
        if (base == first && first < last) st |= 1u;
        if (fail_mask) st |= 2u;
    }
    if (lane == 0) {
        if (kSize)
            out_off[s] = at;
        else
            state[s] = static_cast<uint8_t>(st);
    }
}

// Exclusive scan of n 64-bit values in place by one CTA; total written behind them (values[n]).
__global__ void __launch_bounds__(1024) wire_scan_kernel(unsigned long long* __restrict__ values, const size_t n) {
    __shared__ unsigned long long carry;
    __shared__ unsigned long long warp_tot[32];
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (size_t base = 0; base < n; base += 1024) {
        const size_t i = base + threadIdx.x;
        const unsigned long long v = i < n ? values[i] : 0ull;
        unsigned long long incl = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned long long t = __shfl_up_sync(kFull, incl, o);
            if ((threadIdx.x & 31) >= static_cast<unsigned>(o)) incl += t;
        }
        if ((threadIdx.x & 31) == 31) warp_tot[threadIdx.x >> 5] = incl;
        __syncthreads();
        unsigned long long before = carry;
        for (unsigned w = 0; w < (threadIdx.x >> 5); ++w) before += warp_tot[w];
        if (i < n) values[i] = before + incl - v;
        __syncthreads();
        if (threadIdx.x == 1023) carry = before + incl;
        __syncthreads();
    }
    if (threadIdx.x == 0) values[n] = carry;
}

// ------------------------------------------------------------ host: type definitions
// wireType{StructT: &structType{CommonType{Name, Id}, Field: []fieldType{{Name, Id}...}}} as the
// message gob sends before the first value of a struct type.  [encoding/gob type.go]
void host_uint(std::string& b, uint64_t u) {
    if (u < 128) {
        b.push_back(static_cast<char>(u));
        return;
    }
    int nb = 0;
    for (uint64_t t = u; t; t >>= 8) ++nb;
    b.push_back(static_cast<char>(-nb));
    for (int k = nb - 1; k >= 0; --k) b.push_back(static_cast<char>(u >> (8 * k)));
}
void host_int(std::string& b, int64_t i) {
    host_uint(b, i < 0 ? ((~static_cast<uint64_t>(i)) << 1) | 1ull : static_cast<uint64_t>(i) << 1);
}
void host_string(std::string& b, const char* s) {
    const std::string v(s);
    host_uint(b, v.size());
    b += v;
}
std::string struct_typedef(int id, const char* name, int nfields, const char* const* fnames, const int* fids) {
    std::string b;
    host_int(b, -id);   // a negative id introduces a definition
    b.push_back(3);     // wireType: delta to StructT
    b.push_back(1);     //   structType: delta to CommonType
    b.push_back(1);     //     CommonType.Name
    host_string(b, name);
    b.push_back(1);     //     CommonType.Id
    host_int(b, id);
    b.push_back(0);
    if (nfields) {
        b.push_back(1);  //   structType.Field
        host_uint(b, static_cast<uint64_t>(nfields));
        for (int i = 0; i < nfields; ++i) {
            b.push_back(1);
            host_string(b, fnames[i]);
            b.push_back(1);
            host_int(b, fids[i]);
            b.push_back(0);
        }
    }
    b.push_back(0);
    b.push_back(0);
    std::string msg;
    host_uint(msg, b.size());
    return msg + b;
}

}  // namespace

// The two definitions a server connection may have to send: rpc.Response and the empty struct
// that stands for the reply behind an error.  gob's built-in ids: string 6, uint 3.
void wire_server_typedefs(uint32_t response_id, uint32_t empty_id, std::string* response_def, std::string* empty_def) {
    const char* fn[] = {"ServiceMethod", "Seq", "Error"};
    const int fi[] = {6, 3, 6};
    *response_def = struct_typedef(static_cast<int>(response_id), "Response", 3, fn, fi);
    *empty_def = struct_typedef(static_cast<int>(empty_id), "", 0, nullptr, nullptr);
}

cudaError_t launch_wire_count(const uint8_t* bytes, const unsigned long long* stream_off, size_t n_streams,
                              unsigned long long* call_off, cudaStream_t stream) {
    if (n_streams) {
        const unsigned blocks = static_cast<unsigned>((n_streams + kWireThreads / 32 - 1) / (kWireThreads / 32));
        wire_decode_kernel<true><<<blocks, kWireThreads, 0, stream>>>(bytes, stream_off, n_streams, nullptr, call_off,
                                                                      nullptr, nullptr);
    }
    wire_scan_kernel<<<1, 1024, 0, stream>>>(call_off, n_streams);
    return cudaGetLastError();
}

cudaError_t launch_wire_decode(const uint8_t* bytes, const unsigned long long* stream_off, size_t n_streams,
                               const ptype_wire_stream* streams, unsigned long long* call_off, ptype_req* reqs,
                               ptype_wire_call* calls, cudaStream_t stream) {
    if (n_streams == 0) return cudaSuccess;
    const unsigned blocks = static_cast<unsigned>((n_streams + kWireThreads / 32 - 1) / (kWireThreads / 32));
    wire_decode_kernel<false><<<blocks, kWireThreads, 0, stream>>>(bytes, stream_off, n_streams, streams, call_off, reqs,
                                                                   calls);
    return cudaGetLastError();
}

cudaError_t launch_wire_encode(bool size_only, const uint8_t* req_bytes, const unsigned long long* call_off,
                               size_t n_streams, const ptype_wire_call* calls, const ptype_reply* replies,
                               const uint8_t* defs, uint32_t response_def_len, uint32_t empty_def_len,
                               uint32_t response_id, uint32_t empty_id, uint8_t* state, unsigned long long* out_off,
                               uint8_t* out, cudaStream_t stream) {
    const WireTypes types = {defs, response_def_len, empty_def_len, response_id, empty_id};
    const unsigned blocks = static_cast<unsigned>((n_streams + kWireThreads / 32 - 1) / (kWireThreads / 32));
    if (size_only) {
        if (n_streams)
            wire_encode_kernel<true><<<blocks, kWireThreads, 0, stream>>>(req_bytes, call_off, n_streams, calls, replies,
                                                                          types, state, out_off, out);
        wire_scan_kernel<<<1, 1024, 0, stream>>>(out_off, n_streams);
    } else if (n_streams) {
        wire_encode_kernel<false><<<blocks, kWireThreads, 0, stream>>>(req_bytes, call_off, n_streams, calls, replies,
                                                                       types, state, out_off, out);
    }
    return cudaGetLastError();
}

}  // namespace ptype
