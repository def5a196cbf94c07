// cluster.hpp — C++ mirror of the part of edegens/ptype's `package cluster` that sits on
// the batched-dispatch path, written against the C ABI only (include/ptype_b200.h).
//
// The reference is Go; this image has no Go toolchain, so the host side above the C ABI is
// mirrored here in C++ with the reference's names, argument meaning and error behaviour:
//
//   cluster::Node, cluster::Registry          cluster/registry.go:17-26
//   cluster::MemRegistry                      stands in for etcdRegistry (registry.go:28-170):
//                                             same keys, JSON values, range order, watch rule
//   cluster::ConnConfig, DefaultConnConfig    cluster/rpc.go:19-38
//   cluster::ErrNoClientAvailable             cluster/rpc.go:15-17
//   cluster::Dispatcher                       new: owns the ptype_ctx and the whole-registry
//                                             snapshot image (what Services() returns)
//   cluster::Client{Call,CallBatch,Close,ConnectionErrs,getSelectedNodes}
//                                             cluster/rpc.go:40-124,126-297
//
// What differs, on purpose: connections are not dialled (connectToNodes, rpc.go:272-285, has
// no device analogue), and withRetry retries at most cfg.Retries times (the reference loops
// forever when Retries > 0 and a call keeps failing, rpc.go:107-116).
#pragma once

#include <algorithm>
#include <array>
#include <atomic>
#include <cctype>
#include <chrono>
#include <condition_variable>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <functional>
#include <map>
#include <memory>
#include <mutex>
#include <optional>
#include <string>
#include <thread>
#include <utility>
#include <vector>

#include "../../include/ptype_b200.h"

namespace cluster {

// ----------------------------------------------------------------------------- errors
// Go's `error`: empty message = nil.
struct Error {
    std::string msg;
    Error() = default;
    explicit Error(std::string m) : msg(std::move(m)) {}
    explicit operator bool() const { return !msg.empty(); }  // true = there IS an error
    const std::string& Error_() const { return msg; }
    bool Is(const Error& other) const { return msg == other.msg; }
};
inline const Error& ErrNoClientAvailable() {  // cluster/rpc.go:15-17
    static const Error e("no client nodes available");
    return e;
}

// ------------------------------------------------------------------------------- Node
struct Node {  // cluster/registry.go:23-26
    std::string Address;
    int Port = 0;
    bool operator==(const Node& o) const { return Address == o.Address && Port == o.Port; }
    bool operator!=(const Node& o) const { return !(*this == o); }
};

namespace detail {

// json.Marshal(Node) (registry.go:52-53): {"address":"...","port":N}
inline std::string MarshalNode(const Node& n) {
    std::string out = "{\"address\":\"";
    for (unsigned char c : n.Address) {
        char buf[8];
        if (c == '"' || c == '\\') {
            out.push_back('\\');
            out.push_back(static_cast<char>(c));
        } else if (c == '\n') {
            out += "\\n";
        } else if (c == '\r') {
            out += "\\r";
        } else if (c == '\t') {
            out += "\\t";
        } else if (c < 0x20 || c == '<' || c == '>' || c == '&') {
            std::snprintf(buf, sizeof buf, "\\u%04x", c);
            out += buf;
        } else {
            out.push_back(static_cast<char>(c));
        }
    }
    out += "\",\"port\":" + std::to_string(n.Port) + "}";
    return out;
}

// json.Unmarshal into an existing Node: absent fields keep their value (registry.go:99,102).
// Accepts the JSON the reference writes and its tests insert by hand; false = error.
inline bool UnmarshalNode(const std::string& js, Node* n) {
    size_t i = 0;
    auto ws = [&] { while (i < js.size() && (js[i] == ' ' || js[i] == '\t' || js[i] == '\n' || js[i] == '\r')) ++i; };
    auto str = [&](std::string* out) {
        if (i >= js.size() || js[i] != '"') return false;
        ++i;
        out->clear();
        while (i < js.size() && js[i] != '"') {
            char c = js[i++];
            if (c == '\\') {
                if (i >= js.size()) return false;
                char e = js[i++];
                switch (e) {
                    case 'n': out->push_back('\n'); break;
                    case 'r': out->push_back('\r'); break;
                    case 't': out->push_back('\t'); break;
                    case 'b': out->push_back('\b'); break;
                    case 'f': out->push_back('\f'); break;
                    case 'u': {
                        if (i + 4 > js.size()) return false;
                        unsigned cp = static_cast<unsigned>(std::stoul(js.substr(i, 4), nullptr, 16));
                        i += 4;
                        if (cp < 0x80) {
                            out->push_back(static_cast<char>(cp));
                        } else if (cp < 0x800) {
                            out->push_back(static_cast<char>(0xC0 | (cp >> 6)));
                            out->push_back(static_cast<char>(0x80 | (cp & 0x3F)));
                        } else {
                            out->push_back(static_cast<char>(0xE0 | (cp >> 12)));
                            out->push_back(static_cast<char>(0x80 | ((cp >> 6) & 0x3F)));
                            out->push_back(static_cast<char>(0x80 | (cp & 0x3F)));
                        }
                        break;
                    }
                    default: out->push_back(e);
                }
            } else {
                out->push_back(c);
            }
        }
        if (i >= js.size()) return false;
        ++i;
        return true;
    };
    ws();
    if (i >= js.size() || js[i] != '{') return false;
    ++i;
    ws();
    if (i < js.size() && js[i] == '}') return true;
    for (;;) {
        std::string key;
        ws();
        if (!str(&key)) return false;
        for (char& c : key) c = static_cast<char>(std::tolower(static_cast<unsigned char>(c)));
        ws();
        if (i >= js.size() || js[i] != ':') return false;
        ++i;
        ws();
        if (key == "address") {
            std::string v;
            if (!str(&v)) return false;
            n->Address = v;
        } else if (key == "port") {
            size_t j = i;
            if (j < js.size() && js[j] == '-') ++j;
            size_t d0 = j;
            while (j < js.size() && js[j] >= '0' && js[j] <= '9') ++j;
            if (j == d0 || (j < js.size() && (js[j] == '.' || js[j] == 'e' || js[j] == 'E'))) return false;
            n->Port = std::stoi(js.substr(i, j - i));
            i = j;
        } else {  // unknown field: skip a scalar or string
            if (i < js.size() && js[i] == '"') {
                std::string skip;
                if (!str(&skip)) return false;
            } else {
                while (i < js.size() && js[i] != ',' && js[i] != '}') ++i;
            }
        }
        ws();
        if (i < js.size() && js[i] == ',') {
            ++i;
            continue;
        }
        if (i < js.size() && js[i] == '}') return true;
        return false;
    }
}

// etcdKey (registry.go:168-170) for names without '/', '.' or '..' elements.
inline std::string EtcdKey(std::initializer_list<std::string> elems) {
    std::string k;
    for (const std::string& e : elems) {
        if (e.empty()) continue;
        if (!k.empty()) k.push_back('/');
        k += e;
    }
    return k + "/";
}

}  // namespace detail

// ---------------------------------------------------------------------------- channels
// `chan []Node` of WatchService: blocking receive with timeout, close by the sender.
class NodeChan {
   public:
    void Send(std::vector<Node> v) {
        std::lock_guard<std::mutex> lk(mu_);
        if (closed_) return;
        q_.push_back(std::move(v));
        cv_.notify_all();
    }
    void Close() {
        std::lock_guard<std::mutex> lk(mu_);
        closed_ = true;
        cv_.notify_all();
    }
    // nullopt on timeout; ok=false once closed and drained (Go: v, ok := <-ch)
    std::optional<std::vector<Node>> Recv(std::chrono::milliseconds timeout, bool* ok) {
        std::unique_lock<std::mutex> lk(mu_);
        *ok = true;
        if (!cv_.wait_for(lk, timeout, [&] { return !q_.empty() || closed_; })) return std::nullopt;
        if (q_.empty()) {
            *ok = false;
            return std::nullopt;
        }
        std::vector<Node> v = std::move(q_.front());
        q_.pop_front();
        return v;
    }

   private:
    std::mutex mu_;
    std::condition_variable cv_;
    std::deque<std::vector<Node>> q_;
    bool closed_ = false;
};

// ---------------------------------------------------------------------------- Registry
class Registry {  // cluster/registry.go:17-21
   public:
    virtual ~Registry() = default;
    virtual Error Register(const std::string& serviceName, const std::string& nodeName,
                           const std::string& host, int port) = 0;
    virtual Error Services(std::map<std::string, std::vector<Node>>* out) = 0;
    // Sends the current list first, then the full list after every change under the
    // service's prefix (registry.go:119-150); Cancel closes the channel.
    virtual std::shared_ptr<NodeChan> WatchService(const std::string& serviceName) = 0;
    virtual void CancelWatch(const std::shared_ptr<NodeChan>& ch) = 0;
};

// In-memory stand-in for etcdRegistry: an ordered KV with the reference's key layout.
class MemRegistry : public Registry {
   public:
    Error Register(const std::string& serviceName, const std::string& nodeName, const std::string& host,
                   int port) override {  // registry.go:51-86 (lease/keep-alive: no stored effect)
        Put(detail::EtcdKey({"services", serviceName, nodeName}), detail::MarshalNode(Node{host, port}));
        return Error();
    }
    void Put(const std::string& key, const std::string& value) {
        std::vector<std::pair<std::shared_ptr<NodeChan>, std::vector<Node>>> notify;
        {
            std::lock_guard<std::mutex> lk(mu_);
            kv_[key] = value;
            CollectLocked(key, &notify);
        }
        for (auto& n : notify) n.first->Send(std::move(n.second));
    }
    int Delete(const std::string& key) {
        std::vector<std::pair<std::shared_ptr<NodeChan>, std::vector<Node>>> notify;
        int deleted;
        {
            std::lock_guard<std::mutex> lk(mu_);
            deleted = static_cast<int>(kv_.erase(key));
            if (deleted) CollectLocked(key, &notify);
        }
        for (auto& n : notify) n.first->Send(std::move(n.second));
        return deleted;
    }
    // registry.go:93-117: ONE range over the prefix "services", ONE reused Node variable.
    Error Services(std::map<std::string, std::vector<Node>>* out) override {
        std::lock_guard<std::mutex> lk(mu_);
        out->clear();
        Node node;
        for (auto it = kv_.lower_bound("services"); it != kv_.end() && it->first.compare(0, 8, "services") == 0; ++it) {
            if (!detail::UnmarshalNode(it->second, &node)) return Error("failed to unmarshal services nodes");
            size_t s1 = it->first.find('/');
            if (s1 == std::string::npos) continue;
            size_t s2 = it->first.find('/', s1 + 1);
            (*out)[it->first.substr(s1 + 1, s2 == std::string::npos ? std::string::npos : s2 - s1 - 1)].push_back(node);
        }
        return Error();
    }
    // registry.go:152-166
    Error nodes(const std::string& serviceName, std::vector<Node>* out) {
        std::lock_guard<std::mutex> lk(mu_);
        return NodesLocked(serviceName, out);
    }
    std::shared_ptr<NodeChan> WatchService(const std::string& serviceName) override {
        auto ch = std::make_shared<NodeChan>();
        std::vector<Node> now;
        {
            std::lock_guard<std::mutex> lk(mu_);
            NodesLocked(serviceName, &now);
            watchers_.push_back({serviceName, ch});
        }
        ch->Send(std::move(now));
        return ch;
    }
    void CancelWatch(const std::shared_ptr<NodeChan>& ch) override {
        {
            std::lock_guard<std::mutex> lk(mu_);
            for (auto it = watchers_.begin(); it != watchers_.end();)
                it = it->second == ch ? watchers_.erase(it) : it + 1;
        }
        ch->Close();
    }

   private:
    Error NodesLocked(const std::string& serviceName, std::vector<Node>* out) {
        const std::string prefix = detail::EtcdKey({"services", serviceName});  // trailing '/' matters
        out->clear();
        for (auto it = kv_.lower_bound(prefix); it != kv_.end() && it->first.compare(0, prefix.size(), prefix) == 0; ++it) {
            Node n;
            if (!detail::UnmarshalNode(it->second, &n)) return Error("failed to unmarshal services nodes");
            out->push_back(n);
        }
        return Error();
    }
    void CollectLocked(const std::string& key, std::vector<std::pair<std::shared_ptr<NodeChan>, std::vector<Node>>>* notify) {
        for (auto& w : watchers_) {
            const std::string prefix = detail::EtcdKey({"services", w.first});
            if (key.compare(0, prefix.size(), prefix) != 0) continue;
            std::vector<Node> now;
            NodesLocked(w.first, &now);
            notify->push_back({w.second, std::move(now)});
        }
    }
    std::mutex mu_;
    std::map<std::string, std::string> kv_;  // byte-ordered like the etcd range
    std::vector<std::pair<std::string, std::shared_ptr<NodeChan>>> watchers_;
};

// ----------------------------------------------------------------------------- KVStore
// The slice of cluster/store.go:38-62 that snapshot publication needs: Put, and
// Get with the prefix option, over an in-memory ordered map standing in for etcd (not in this
// image).  Keys are filepath.Join("store", key) as in the reference.  The rest of KVStore stays in
// Go: it is not on the dispatch path.
inline const Error& ErrNoKey() {  // store.go:14-16
    static const Error e("Key could not be found");
    return e;
}
class KVStore {
   public:
    static std::string StoreKey(const std::string& key) {  // filepath.Join drops a trailing '/'
        std::string k = "store/" + key;
        while (k.size() > 1 && k.back() == '/') k.pop_back();
        return k;
    }
    Error Get(const std::string& key, std::vector<std::string>* out, bool withPrefix = false) {  // store.go:38-54
        std::lock_guard<std::mutex> lk(mu_);
        out->clear();
        const std::string k = StoreKey(key);
        if (withPrefix) {
            for (auto it = kv_.lower_bound(k); it != kv_.end() && it->first.compare(0, k.size(), k) == 0; ++it)
                out->push_back(it->second);
        } else {
            auto it = kv_.find(k);
            if (it != kv_.end()) out->push_back(it->second);
        }
        return out->empty() ? ErrNoKey() : Error();
    }
    Error Put(const std::string& key, const std::string& value) {  // store.go:56-62
        std::lock_guard<std::mutex> lk(mu_);
        kv_[StoreKey(key)] = value;
        return Error();
    }
    // A registry image under store/registry_image/<epoch, zero padded>: one host builds and
    // publishes, every host of the deployment installs the same epoch.
    Error PublishImage(uint64_t epoch, const void* blob, size_t bytes) {
        char name[64];
        std::snprintf(name, sizeof name, "registry_image/%020llu", static_cast<unsigned long long>(epoch));
        return Put(name, std::string(static_cast<const char*>(blob), bytes));
    }
    Error LatestImage(std::string* blob) {
        std::vector<std::string> all;
        if (Error e = Get("registry_image/", &all, /*withPrefix=*/true)) return e;
        *blob = all.back();  // keys sort by epoch
        return Error();
    }

   private:
    std::mutex mu_;
    std::map<std::string, std::string> kv_;
};

// ----------------------------------------------------------------------------- optimus
// example/optimus/coordinator/coordinator.go:67-99: the call stream one target generates and
// the rule that picks the answer.
namespace optimus {
struct Args {  // example/optimus/prime.go:7-11
    int64_t Min, Max, Target;
};
inline std::vector<Args> SplitWork(int64_t target) {  // coordinator.go:67-73
    std::vector<Args> out;
    int64_t min = 2;
    const int64_t limit = static_cast<int64_t>(static_cast<uint64_t>(target) + 10u);
    for (int64_t i = 10; i < limit; i += 10) {
        out.push_back(Args{min, i, target});
        min = i;
    }
    return out;
}
// watchReplies (coordinator.go:91-99) on replies in arrival order: the first one that is not
// the target, else the target.
inline int64_t WatchReplies(int64_t target, const std::vector<int64_t>& replies) {
    for (int64_t r : replies)
        if (r != target) return r;
    return target;
}
}  // namespace optimus

// -------------------------------------------------------------------------- ConnConfig
struct ConnConfig {  // cluster/rpc.go:19-31
    int MaxConnections = 3;
    std::chrono::milliseconds InitialNodeTimeout{5000};
    std::chrono::milliseconds DebounceTime{3000};
    int Retries = 2;
};
inline const ConnConfig& DefaultConnConfig() {  // cluster/rpc.go:33-38
    static const ConnConfig c;
    return c;
}

// -------------------------------------------------------------------------- Dispatcher
// Owns the device context and the registry snapshot shared by every Client of this host:
// per service, the last node list its Client accepted.  Rebuilding and publishing the
// image is the device form of handleNewNodes' swap (rpc.go:226-244).
class Dispatcher {
   public:
    Dispatcher(std::string localAddr, int maxConnections, int device = 0)
        : localAddr_(std::move(localAddr)), maxConnections_(maxConnections) {
        int rc = ptype_b200_open(&device, 1, &ctx_);
        if (rc != PTYPE_OK) err_ = Error(std::string("ptype_b200_open: ") + ptype_b200_strerror(rc) + ": " + ptype_b200_last_error());
    }
    ~Dispatcher() {
        if (ctx_) ptype_b200_close(ctx_);
        if (image_) ptype_b200_image_free(image_);
    }
    Dispatcher(const Dispatcher&) = delete;
    Dispatcher& operator=(const Dispatcher&) = delete;
    const Error& Err() const { return err_; }
    ptype_ctx* ctx() { return ctx_; }
    int MaxConnections() const { return maxConnections_; }

    // rpc.Register(new(T)) on the nodes of a service (server.go:16-17).
    void Expose(const std::string& service, const std::string& receiver) {
        std::lock_guard<std::mutex> lk(mu_);
        exposed_[service].push_back(receiver);
    }
    // Install `nodes` as the service's list and publish a new epoch.
    Error SetNodes(const std::string& service, const std::vector<Node>& nodes) {
        std::lock_guard<std::mutex> lk(mu_);
        lists_[service] = nodes;
        return PublishLocked();
    }
    uint64_t ServiceKey(const std::string& service) {
        std::lock_guard<std::mutex> lk(mu_);
        return ptype_b200_service_key(service.c_str(), salt_);
    }
    // getSelectedNodes() (rpc.go:293-297)
    std::vector<Node> SelectedNodes(const std::string& service) {
        std::lock_guard<std::mutex> lk(mu_);
        std::vector<Node> out;
        if (!image_) return out;
        uint32_t k = 0;
        std::vector<uint32_t> ids(4096);
        if (ptype_b200_image_selected(image_, service.c_str(), ids.data(), ids.size(), &k) != PTYPE_OK) return out;
        if (k > ids.size()) {
            ids.resize(k);
            ptype_b200_image_selected(image_, service.c_str(), ids.data(), ids.size(), &k);
        }
        for (uint32_t i = 0; i < k; ++i) out.push_back(NodeOfLocked(ids[i]));
        return out;
    }
    Node NodeOf(uint32_t id) {
        std::lock_guard<std::mutex> lk(mu_);
        return NodeOfLocked(id);
    }
    uint64_t Epoch() {
        std::lock_guard<std::mutex> lk(mu_);
        return epoch_;
    }

    // Snapshot publication through the KV store.  PublishImage stores the image
    // this Dispatcher last built; LoadLatestImage, on another host (another Dispatcher), installs
    // the newest published epoch on its GPU: both then answer from the same bytes.
    Error PublishImage(KVStore* store) {
        std::lock_guard<std::mutex> lk(mu_);
        if (!image_) return Error("ptype_b200: no image built yet");
        const ptype_image_header* h = static_cast<const ptype_image_header*>(image_);
        return store->PublishImage(h->epoch, image_, h->total_bytes);
    }
    Error LoadLatestImage(KVStore* store) {
        std::string blob;
        if (Error e = store->LatestImage(&blob)) return e;
        std::lock_guard<std::mutex> lk(mu_);
        int rc = ptype_b200_image_validate(blob.data(), blob.size());
        if (rc == PTYPE_OK) rc = ptype_b200_publish_image(ctx_, blob.data(), blob.size());
        if (rc != PTYPE_OK) return Error(std::string("install image: ") + ptype_b200_strerror(rc) + ": " + ptype_b200_last_error());
        void* copy = std::malloc(blob.size());  // image_free() is free()
        if (!copy) return Error("out of memory");
        std::memcpy(copy, blob.data(), blob.size());
        if (image_) ptype_b200_image_free(image_);
        image_ = copy;
        const ptype_image_header* h = static_cast<const ptype_image_header*>(image_);
        epoch_ = h->epoch;
        salt_ = h->key_salt;
        return Error();
    }

    // One batch for MANY services: what a process holding one Client per service does when it
    // issues reqs[0], reqs[1], ... in order, each through the Client of its own service.  Every
    // Client draws its sequence number from its own counter (rpc.go:129,181) and only when its
    // client list is not empty (rpc.go:178-180); here those counters live on the device and the
    // seq field of the records is filled there (ptype_b200_assign_seq), so callers set svc_key,
    // method, recv and the arguments only.
    Error CallMixed(std::vector<ptype_req>& reqs, std::vector<ptype_reply>* replies) {
        replies->assign(reqs.size(), ptype_reply{});
        if (reqs.empty()) return Error();
        const size_t n = reqs.size();
        void *d_req = nullptr, *d_rep = nullptr, *d_seq = nullptr;
        auto fail = [&](const char* what, int rc) {
            if (d_req) ptype_b200_device_free(ctx_, 0, d_req);
            if (d_rep) ptype_b200_device_free(ctx_, 0, d_rep);
            if (d_seq) ptype_b200_device_free(ctx_, 0, d_seq);
            return Error(std::string(what) + ": " + ptype_b200_strerror(rc) + ": " + ptype_b200_last_error());
        };
        int rc = ptype_b200_device_alloc(ctx_, 0, n * sizeof(ptype_req), &d_req);
        if (rc != PTYPE_OK) return fail("device_alloc", rc);
        rc = ptype_b200_device_alloc(ctx_, 0, n * sizeof(ptype_reply), &d_rep);
        if (rc != PTYPE_OK) return fail("device_alloc", rc);
        rc = ptype_b200_device_alloc(ctx_, 0, n * sizeof(uint64_t), &d_seq);
        if (rc != PTYPE_OK) return fail("device_alloc", rc);
        rc = ptype_b200_memcpy_h2d(ctx_, 0, d_req, reqs.data(), n * sizeof(ptype_req));
        if (rc != PTYPE_OK) return fail("memcpy_h2d", rc);
        // the numbers go to an array beside the records and the dispatch reads them there: storing
        // them into the 48-byte records would cost the GPU a read and a write of a sector per call
        rc = ptype_b200_assign_seq_array(ctx_, 0, static_cast<const ptype_req*>(d_req), n,
                                         static_cast<uint64_t*>(d_seq), PTYPE_F_ASYNC);
        if (rc != PTYPE_OK) return fail("assign_seq_array", rc);
        rc = ptype_b200_dispatch_seq(ctx_, 0, static_cast<const ptype_req*>(d_req), static_cast<const uint64_t*>(d_seq),
                                     n, static_cast<ptype_reply*>(d_rep), 0);
        if (rc != PTYPE_OK) return fail("dispatch_seq", rc);
        rc = ptype_b200_memcpy_d2h(ctx_, 0, replies->data(), d_rep, n * sizeof(ptype_reply));
        if (rc != PTYPE_OK) return fail("memcpy_d2h", rc);
        std::vector<uint64_t> seq(n);
        rc = ptype_b200_memcpy_d2h(ctx_, 0, seq.data(), d_seq, n * sizeof(uint64_t));  // the numbers drawn
        if (rc != PTYPE_OK) return fail("memcpy_d2h", rc);
        for (size_t j = 0; j < n; ++j) reqs[j].seq = seq[j];
        ptype_b200_device_free(ctx_, 0, d_req);
        ptype_b200_device_free(ctx_, 0, d_rep);
        ptype_b200_device_free(ctx_, 0, d_seq);
        return Error();
    }
    // The device-side counter of a service's Client (the last value drawn).
    uint64_t Seq(const std::string& service) {
        uint64_t v = 0;
        ptype_b200_seq_get(ctx_, 0, service.c_str(), &v);
        return v;
    }

   private:
    Node NodeOfLocked(uint32_t id) {
        char addr[256];
        int port = 0;
        Node n;
        if (image_ && ptype_b200_image_node(image_, id, addr, sizeof addr, &port, nullptr, 0, nullptr, 0) == PTYPE_OK) {
            n.Address = addr;
            n.Port = port;
        }
        return n;
    }
    Error PublishLocked() {
        ptype_builder* b = nullptr;
        int rc = ptype_b200_builder_new(localAddr_.c_str(), maxConnections_, &b);
        if (rc != PTYPE_OK) return Error(std::string("builder_new: ") + ptype_b200_strerror(rc));
        for (const auto& kv : lists_) {
            ptype_b200_builder_declare(b, kv.first.c_str());
            char name[32];
            for (size_t i = 0; i < kv.second.size(); ++i) {
                std::snprintf(name, sizeof name, "%08zu", i);  // keeps the watch's order under byte sort
                rc = ptype_b200_builder_register(b, kv.first.c_str(), name, kv.second[i].Address.c_str(), kv.second[i].Port);
                if (rc != PTYPE_OK) break;
            }
            auto ex = exposed_.find(kv.first);
            if (ex != exposed_.end())
                for (const std::string& r : ex->second) ptype_b200_builder_expose(b, kv.first.c_str(), r.c_str());
        }
        void* blob = nullptr;
        size_t bytes = 0;
        if (rc == PTYPE_OK) rc = ptype_b200_builder_finish(b, epoch_ + 1, &blob, &bytes);
        ptype_b200_builder_free(b);
        if (rc != PTYPE_OK) return Error(std::string("image build: ") + ptype_b200_strerror(rc));
        rc = ptype_b200_publish_image(ctx_, blob, bytes);
        if (rc != PTYPE_OK) {
            ptype_b200_image_free(blob);
            return Error(std::string("publish_image: ") + ptype_b200_strerror(rc) + ": " + ptype_b200_last_error());
        }
        if (image_) ptype_b200_image_free(image_);
        image_ = blob;
        ++epoch_;
        salt_ = static_cast<const ptype_image_header*>(blob)->key_salt;
        return Error();
    }
    std::string localAddr_;
    int maxConnections_;
    ptype_ctx* ctx_ = nullptr;
    Error err_;
    std::mutex mu_;
    std::map<std::string, std::vector<Node>> lists_;
    std::map<std::string, std::vector<std::string>> exposed_;
    void* image_ = nullptr;
    uint64_t epoch_ = 0, salt_ = 0;
};

// ------------------------------------------------------------------------------ Client
struct BatchReply {  // mirrors ptype_reply
    int64_t Value = 0;
    uint32_t NodeId = PTYPE_NO_NODE;
    uint32_t Status = 0;
};

// What Client.Call would have returned for a per-call status.
inline Error StatusError(uint32_t status, const std::string& serviceMethod) {
    switch (status) {
        case PTYPE_ST_OK: return Error();
        case PTYPE_ST_NO_SERVICE:
        case PTYPE_ST_NO_CLIENT: return ErrNoClientAvailable();                                     // rpc.go:62-64
        case PTYPE_ST_ILL_FORMED: return Error("rpc: service/method request ill-formed: " + serviceMethod);  // [net/rpc]
        case PTYPE_ST_CANT_FIND_SERVICE: return Error("rpc: can't find service " + serviceMethod);
        case PTYPE_ST_CANT_FIND_METHOD: return Error("rpc: can't find method " + serviceMethod);
        case PTYPE_ST_BAD_INDEX: return Error("runtime error: index out of range");
        case PTYPE_ST_RANGE_LIMIT: return Error("ptype_b200: Prime.Check range beyond the dispatch limit");
        default: return Error("ptype_b200: unknown status");
    }
}

class Client {  // cluster/rpc.go:40-43,126-139
   public:
    // newClient + newConnectionBalancer (rpc.go:45-57,141-170): waits InitialNodeTimeout for the
    // first list ("no initial nodes provided for <svc>"), applies it, then watches with debounce.
    static Error New(Dispatcher* gpu, const std::string& serviceName, Registry* r, const ConnConfig* cfg,
                     std::unique_ptr<Client>* out) {
        std::unique_ptr<Client> c(new Client(gpu, serviceName, r, cfg ? *cfg : DefaultConnConfig()));
        if (c->cfg_.MaxConnections != gpu->MaxConnections())
            return Error("ptype_b200: ConnConfig.MaxConnections must equal the Dispatcher's (one selection rule per image)");
        c->nodesChan_ = r->WatchService(serviceName);
        bool ok = true;
        auto initial = c->nodesChan_->Recv(c->cfg_.InitialNodeTimeout, &ok);
        if (!initial) {
            r->CancelWatch(c->nodesChan_);
            return Error("no initial nodes provided for " + serviceName);  // rpc.go:159
        }
        if (Error e = c->handleNewNodes(*initial)) {
            r->CancelWatch(c->nodesChan_);
            return e;
        }
        c->watcher_ = std::thread([p = c.get()] { p->watchForNewNodes(); });
        *out = std::move(c);
        return Error();
    }
    ~Client() { Close(); }

    // rpc.go:59-67.  args = calculator.Args{A,B} or prime.Args{Min,Max,Target}.
    Error Call(const std::string& serviceMethod, const int64_t (&args)[3], int64_t* reply) {
        BatchReply rep;
        Error e = CallBatch(serviceMethod, &args, 1, &rep);
        if (e) return e;
        if (Error se = StatusError(rep.Status, serviceMethod)) return se;  // reply untouched on error
        *reply = rep.Value;
        return Error();
    }

    // ------------------------------------------------------------------ batches
    // Page-locked buffers for one Client's batches, allocated once and reused from call to call.
    // The GPU reads Args()/Sel() and writes Values() (or Replies()) across PCIe itself, so what
    // the caller puts there is what travels: 16-25 bytes per call in, 8 out, no 48-byte request
    // records and no staging copy.  (The Go side does the same with slices over C-owned pinned
    // memory, which is also what the cgo pointer rules want.)
    class Batch {
       public:
        Batch(size_t capacity, int argsPerCall) : cap_(capacity ? capacity : 1), apc_(argsPerCall) {
            args_ = static_cast<int64_t*>(ptype_b200_alloc_pinned(cap_ * apc_ * sizeof(int64_t)));
            sel_ = static_cast<uint8_t*>(ptype_b200_alloc_pinned(cap_));
            values_ = static_cast<int64_t*>(ptype_b200_alloc_pinned(cap_ * sizeof(int64_t)));
        }
        ~Batch() {
            ptype_b200_free_pinned(args_);
            ptype_b200_free_pinned(sel_);
            ptype_b200_free_pinned(values_);
            ptype_b200_free_pinned(replies_);
        }
        Batch(const Batch&) = delete;
        Batch& operator=(const Batch&) = delete;
        bool Ok() const { return args_ && sel_ && values_; }
        size_t Capacity() const { return cap_; }
        int ArgsPerCall() const { return apc_; }
        int64_t* Args() { return args_; }      // [n][ArgsPerCall]: calculator.Args{A,B} / prime.Args{Min,Max,Target}
        uint8_t* Sel() { return sel_; }        // [n]: index into the method list of the call (mixed batches)
        int64_t* Values() { return values_; }  // [n]: *reply of every call, 0 for a failed one
        // The calls whose status is not OK after the last attempt: (index in the batch, status).
        const std::vector<ptype_failed>& Failed() const { return failed_; }

       private:
        friend class Client;
        ptype_reply* Replies() {  // 16-byte records (value, node, status): only the classic CallBatch wants them
            if (!replies_) replies_ = static_cast<ptype_reply*>(ptype_b200_alloc_pinned(cap_ * sizeof(ptype_reply)));
            return replies_;
        }
        size_t cap_;
        int apc_;
        int64_t* args_ = nullptr;
        uint8_t* sel_ = nullptr;
        int64_t* values_ = nullptr;
        ptype_reply* replies_ = nullptr;
        std::vector<ptype_failed> failed_, scratch_;
        std::unique_ptr<ptype_failed[]> rawFailed_;  // where the C ABI writes the failed list (capacity cap_)
        std::unique_ptr<Batch> retry_;  // the failed calls of an attempt, re-issued from here
    };
    std::unique_ptr<Batch> NewBatch(size_t capacity, int argsPerCall = 2) {
        return std::unique_ptr<Batch>(new Batch(capacity, argsPerCall));
    }

    // The batch call: n calls whose arguments (and, with more than one method, method selectors)
    // the caller has written into `b`.  Routed exactly as n sequential Call()s: the sequence range
    // is reserved with ONE atomic add (rpc.go:181 batched).  Results: b.Values()[j] and b.Failed().
    // Calls that failed are re-issued with fresh sequence numbers up to cfg.Retries more times
    // (withRetry, rpc.go:107-116, bounded - see the header comment).
    Error CallBatch(const std::vector<std::string>& serviceMethods, Batch& b, size_t n) {
        return CallBatchAttempts(serviceMethods, b, n, cfg_.Retries);
    }
    Error CallBatchAttempts(const std::vector<std::string>& serviceMethods, Batch& b, size_t n, int retries) {
        b.failed_.clear();
        if (n == 0) return Error();
        if (n > b.cap_ || !b.Ok()) return Error("ptype_b200: batch larger than its buffers");
        if (serviceMethods.empty() || serviceMethods.size() > PTYPE_BATCH_METHODS)
            return Error("ptype_b200: a batch names 1 to 4 methods");
        ptype_method_ref refs[PTYPE_BATCH_METHODS];
        for (size_t i = 0; i < serviceMethods.size(); ++i)
            ptype_b200_method_key(serviceMethods[i].c_str(), &refs[i].method, &refs[i].recv);
        const int nm = static_cast<int>(serviceMethods.size());
        const uint64_t key = gpu_->ServiceKey(service_);
        if (Error e = submitValues(key, refs, nm, b, n, &b.failed_)) return e;
        // withRetry: only the calls that failed go round again, each with a fresh sequence number
        for (int attempt = 0; attempt < retries && !b.failed_.empty(); ++attempt) {
            const size_t m = b.failed_.size();
            std::sort(b.failed_.begin(), b.failed_.end(),
                      [](const ptype_failed& x, const ptype_failed& y) { return x.index < y.index; });
            if (!b.retry_ || b.retry_->cap_ < m) b.retry_.reset(new Batch(m, b.apc_));
            Batch& r = *b.retry_;
            if (!r.Ok()) return Error("ptype_b200: out of page-locked memory");
            for (size_t t = 0; t < m; ++t) {
                const size_t j = b.failed_[t].index;
                for (int k = 0; k < b.apc_; ++k) r.args_[t * b.apc_ + k] = b.args_[j * b.apc_ + k];
                r.sel_[t] = b.sel_[j];
            }
            if (Error e = submitValues(key, refs, nm, r, m, &b.scratch_)) return e;
            for (size_t t = 0; t < m; ++t) b.values_[b.failed_[t].index] = r.values_[t];
            for (ptype_failed& f : b.scratch_) f.index = b.failed_[f.index].index;
            b.failed_.swap(b.scratch_);
        }
        return Error();
    }

    // The same through plain arrays, with the routing decision of every call in the reply: args
    // are copied into a pooled Batch, the GPU writes 16-byte reply records, and those are unpacked.
    Error CallBatch(const std::string& serviceMethod, const int64_t (*args)[3], size_t n, BatchReply* replies) {
        return CallBatchAttempts(serviceMethod, args, n, replies, cfg_.Retries);
    }
    Error CallBatchAttempts(const std::string& serviceMethod, const int64_t (*args)[3], size_t n,
                            BatchReply* replies, int retries) {
        if (n == 0) return Error();
        ptype_method_ref ref;
        ptype_b200_method_key(serviceMethod.c_str(), &ref.method, &ref.recv);
        const uint64_t key = gpu_->ServiceKey(service_);
        int apc = 2;
        for (size_t j = 0; j < n && apc == 2; ++j)
            if (args[j][2] != 0) apc = 3;
        std::unique_ptr<Batch> b = acquire(n, apc);
        if (!b->Ok() || !b->Replies()) return Error("ptype_b200: out of page-locked memory");
        std::vector<uint32_t> todo(n);
        for (size_t j = 0; j < n; ++j) todo[j] = static_cast<uint32_t>(j);
        Error err;
        for (int attempt = 0;; ++attempt) {
            const size_t m = todo.size();
            for (size_t t = 0; t < m; ++t)
                for (int k = 0; k < apc; ++k) b->args_[t * apc + k] = args[todo[t]][k];
            const uint64_t base = seq_.fetch_add(m);  // value before the add; the first call sees base+1
            int rc = ptype_b200_call_batch(gpu_->ctx(), 0, key, base, &ref, 1, nullptr, b->args_, apc, m,
                                           b->replies_, 0);
            if (rc != PTYPE_OK) {
                err = Error(std::string("ptype_b200_call_batch: ") + ptype_b200_strerror(rc) + ": " + ptype_b200_last_error());
                break;
            }
            std::vector<uint32_t> next;
            for (size_t t = 0; t < m; ++t) {
                const ptype_reply& o = b->replies_[t];
                BatchReply& out = replies[todo[t]];
                out.Value = o.value;
                out.NodeId = o.node;
                out.Status = o.status;
                if (o.status != PTYPE_ST_OK) next.push_back(todo[t]);
            }
            if (next.empty() || attempt >= retries) break;
            todo.swap(next);
        }
        release(std::move(b));
        return err;
    }

    // rpc.go:69-105.  The asynchronous twin: the call runs on its own thread and the returned
    // object is signalled when done.  Like the reference it makes ONE attempt (rpc.go:91 tests
    // the outer call's Error, which is still nil, so the retry loop never repeats - SURVEY.md
    // D6); on failure the delivered call carries the error and *reply is left untouched
    // (rpc_test.go:262-292).
    struct AsyncCall {
        std::string ServiceMethod;
        Error Err;
        bool Done() {
            std::lock_guard<std::mutex> lk(mu);
            return done;
        }
        void Wait() {
            std::unique_lock<std::mutex> lk(mu);
            cv.wait(lk, [&] { return done; });
        }
        ~AsyncCall() {
            if (worker.joinable()) worker.join();
        }

       private:
        friend class Client;
        friend class CallCoalescer;
        std::mutex mu;
        std::condition_variable cv;
        bool done = false;
        std::thread worker;
    };
    std::shared_ptr<AsyncCall> Go(const std::string& serviceMethod, const int64_t (&args)[3], int64_t* reply) {
        auto call = std::make_shared<AsyncCall>();
        call->ServiceMethod = serviceMethod;
        const int64_t a0 = args[0], a1 = args[1], a2 = args[2];
        AsyncCall* raw = call.get();
        call->worker = std::thread([this, raw, serviceMethod, a0, a1, a2, reply] {
            const int64_t a[1][3] = {{a0, a1, a2}};
            BatchReply rep;
            Error e = CallBatchAttempts(serviceMethod, a, 1, &rep, /*retries=*/0);
            if (!e) e = StatusError(rep.Status, serviceMethod);
            if (!e) *reply = rep.Value;
            std::lock_guard<std::mutex> lk(raw->mu);
            raw->Err = e;
            raw->done = true;
            raw->cv.notify_all();
        });
        return call;
    }

    std::vector<Node> getSelectedNodes() { return gpu_->SelectedNodes(service_); }  // rpc.go:293-297
    uint64_t Seq() const { return seq_.load(); }  // connectionBalancer.seq: the last value drawn
    // Background update failures (rpc.go:122-124,213-218): at most one pending, like chan error, 1.
    std::optional<Error> ConnectionErrs() {
        std::lock_guard<std::mutex> lk(errMu_);
        std::optional<Error> e = pendingErr_;
        pendingErr_.reset();
        return e;
    }
    // Number of updates applied so far (the tests' <-c.connsUpdated).
    uint64_t ConnsUpdated() const { return connsUpdated_.load(); }
    bool WaitConnsUpdated(uint64_t target, std::chrono::milliseconds timeout) {
        auto end = std::chrono::steady_clock::now() + timeout;
        while (connsUpdated_.load() < target) {
            if (std::chrono::steady_clock::now() > end) return false;
            std::this_thread::sleep_for(std::chrono::milliseconds(5));
        }
        return true;
    }
    Error Close() {  // rpc.go:118-120,185-195
        if (closed_.exchange(true)) return Error();
        if (nodesChan_) registry_->CancelWatch(nodesChan_);
        if (watcher_.joinable()) watcher_.join();
        return Error();
    }

   private:
    Client(Dispatcher* gpu, std::string service, Registry* r, ConnConfig cfg)
        : gpu_(gpu), service_(std::move(service)), registry_(r), cfg_(cfg) {}

    // One attempt of a values batch: reserve the sequence range, one C-ABI call, collect the
    // (index, status) entries of the calls that failed.
    Error submitValues(uint64_t key, const ptype_method_ref* refs, int nm, Batch& b, size_t n,
                       std::vector<ptype_failed>* failed) {
        const uint64_t base = seq_.fetch_add(n);  // value before the add; the first call sees base+1
        if (!b.rawFailed_) b.rawFailed_.reset(new ptype_failed[b.cap_]);  // room for every call: never truncated
        uint64_t count = 0;
        int rc = ptype_b200_call_batch_values(gpu_->ctx(), 0, key, base, refs, nm, nm > 1 ? b.sel_ : nullptr, b.args_,
                                              b.apc_, n, b.values_, b.rawFailed_.get(), b.cap_, &count, 0);
        if (rc != PTYPE_OK)
            return Error(std::string("ptype_b200_call_batch_values: ") + ptype_b200_strerror(rc) + ": " +
                         ptype_b200_last_error());
        failed->assign(b.rawFailed_.get(), b.rawFailed_.get() + count);
        return Error();
    }
    // Batches behind the array-based CallBatch: Call() is safe from many threads (rpc.go:59 is
    // goroutine-safe), so each call in flight takes its own buffers from a small pool.
    std::unique_ptr<Batch> acquire(size_t n, int apc) {
        {
            std::lock_guard<std::mutex> lk(poolMu_);
            for (size_t i = 0; i < pool_.size(); ++i) {
                if (pool_[i]->cap_ >= n && pool_[i]->apc_ == apc) {
                    std::unique_ptr<Batch> b = std::move(pool_[i]);
                    pool_.erase(pool_.begin() + static_cast<long>(i));
                    return b;
                }
            }
        }
        size_t cap = 64;
        while (cap < n) cap *= 2;
        return std::unique_ptr<Batch>(new Batch(cap, apc));
    }
    void release(std::unique_ptr<Batch> b) {
        std::lock_guard<std::mutex> lk(poolMu_);
        if (pool_.size() < 8) pool_.push_back(std::move(b));
    }

    // rpc.go:226-244: selectNodes happens inside the image; the swap is the publish.
    Error handleNewNodes(const std::vector<Node>& nodes) {
        if (Error e = gpu_->SetNodes(service_, nodes)) return e;
        connsUpdated_.fetch_add(1);
        return Error();
    }
    // rpc.go:197-224: keep only the latest non-empty list, apply it after DebounceTime of quiet.
    void watchForNewNodes() {
        std::optional<std::vector<Node>> newNodes;
        for (;;) {
            bool ok = true;
            auto got = nodesChan_->Recv(cfg_.DebounceTime, &ok);
            if (!ok) return;  // ctx.Done()
            if (got) {
                if (got->empty()) continue;  // rpc.go:204-206
                newNodes = std::move(got);
                continue;
            }
            if (!newNodes) continue;  // debounce timer fired with nothing pending
            if (Error e = handleNewNodes(*newNodes)) {
                std::lock_guard<std::mutex> lk(errMu_);
                if (!pendingErr_) pendingErr_ = e;
            }
            newNodes.reset();
        }
    }

    Dispatcher* gpu_;
    std::string service_;
    Registry* registry_;
    ConnConfig cfg_;
    std::atomic<uint64_t> seq_{0};  // connectionBalancer.seq (rpc.go:129)
    std::shared_ptr<NodeChan> nodesChan_;
    std::thread watcher_;
    std::atomic<bool> closed_{false};
    std::atomic<uint64_t> connsUpdated_{0};
    std::mutex errMu_;
    std::optional<Error> pendingErr_;
    std::mutex poolMu_;
    std::vector<std::unique_ptr<Batch>> pool_;
};

// -------------------------------------------------------------------------- coalescing
// The dispatch loop a host puts in front of the GPU path: many goroutine-style callers invoke
// Call() concurrently, as they do with cluster.(*Client).Call (rpc.go:59 is goroutine-safe); a
// background thread gathers what arrived within `max_wait` (or `max_batch` calls, whichever comes
// first), submits one CallBatch per method name and hands every caller its own reply or error.
// Sequence numbers are still taken from the Client's single counter, so the calls of a batch
// are routed like consecutive Call()s made in the order they were gathered.
class CallCoalescer {
   public:
    explicit CallCoalescer(Client* client, size_t max_batch = 65536,
                           std::chrono::microseconds max_wait = std::chrono::microseconds(200))
        : client_(client), max_batch_(max_batch), max_wait_(max_wait), worker_([this] { Run(); }) {}
    ~CallCoalescer() {
        {
            std::lock_guard<std::mutex> lk(mu_);
            stop_ = true;
        }
        cv_.notify_all();
        worker_.join();
    }
    // Blocks until the call has been dispatched; same results and errors as Client::Call.
    Error Call(const std::string& serviceMethod, const int64_t (&args)[3], int64_t* reply) {
        Pending p;
        p.args[0] = args[0];
        p.args[1] = args[1];
        p.args[2] = args[2];
        {
            std::unique_lock<std::mutex> lk(mu_);
            queue_[{serviceMethod, false}].push_back(&p);
            ++queued_;
            cv_.notify_all();
            done_cv_.wait(lk, [&] { return p.done; });
        }
        if (p.err) return p.err;
        if (Error se = StatusError(p.rep.Status, serviceMethod)) return se;
        *reply = p.rep.Value;
        return Error();
    }
    // Client.Go without a thread per call: the call joins the next batch and the returned object
    // is signalled by the dispatch loop.  One attempt, error delivered on the call, *reply
    // untouched on failure - the semantics of Client::Go above (rpc.go:69-105).
    std::shared_ptr<Client::AsyncCall> Go(const std::string& serviceMethod, const int64_t (&args)[3], int64_t* reply) {
        auto call = std::make_shared<Client::AsyncCall>();
        call->ServiceMethod = serviceMethod;
        Pending* p = new Pending();
        p->args[0] = args[0];
        p->args[1] = args[1];
        p->args[2] = args[2];
        p->async = call;
        p->reply = reply;
        std::lock_guard<std::mutex> lk(mu_);
        queue_[{serviceMethod, true}].push_back(p);
        ++queued_;
        cv_.notify_all();
        return call;
    }
    uint64_t Batches() const { return batches_.load(); }

   private:
    struct Pending {
        int64_t args[3];
        BatchReply rep;
        Error err;
        bool done = false;
        std::shared_ptr<Client::AsyncCall> async;  // set for Go(): completed by the loop, then freed
        int64_t* reply = nullptr;
    };
    using GroupKey = std::pair<std::string, bool>;  // method, one attempt only (came through Go)
    void Run() {
        for (;;) {
            std::map<GroupKey, std::vector<Pending*>> take;
            {
                std::unique_lock<std::mutex> lk(mu_);
                cv_.wait(lk, [&] { return stop_ || queued_ > 0; });
                if (stop_ && queued_ == 0) return;
                // let the window fill: until max_batch calls are waiting or max_wait has passed
                cv_.wait_for(lk, max_wait_, [&] { return stop_ || queued_ >= max_batch_; });
                take.swap(queue_);
                queued_ = 0;
            }
            for (auto& group : take) {
                std::vector<Pending*>& ps = group.second;
                const std::string& method = group.first.first;
                std::vector<std::array<int64_t, 3>> args(ps.size());
                for (size_t i = 0; i < ps.size(); ++i) args[i] = {ps[i]->args[0], ps[i]->args[1], ps[i]->args[2]};
                std::vector<BatchReply> reps(ps.size());
                const auto a3 = reinterpret_cast<const int64_t(*)[3]>(args.data());
                Error e = group.first.second ? client_->CallBatchAttempts(method, a3, ps.size(), reps.data(), 0)
                                             : client_->CallBatch(method, a3, ps.size(), reps.data());
                batches_.fetch_add(1);
                std::lock_guard<std::mutex> lk(mu_);
                for (size_t i = 0; i < ps.size(); ++i) {
                    Pending* p = ps[i];
                    if (p->async) {
                        Error ce = e ? e : StatusError(reps[i].Status, method);
                        if (!ce) *p->reply = reps[i].Value;
                        {
                            std::lock_guard<std::mutex> alk(p->async->mu);
                            p->async->Err = ce;
                            p->async->done = true;
                        }
                        p->async->cv.notify_all();
                        delete p;
                        continue;
                    }
                    p->rep = reps[i];
                    p->err = e;
                    p->done = true;
                }
            }
            done_cv_.notify_all();
        }
    }
    Client* client_;
    size_t max_batch_;
    std::chrono::microseconds max_wait_;
    std::mutex mu_;
    std::condition_variable cv_, done_cv_;
    std::map<GroupKey, std::vector<Pending*>> queue_;
    size_t queued_ = 0;
    bool stop_ = false;
    std::atomic<uint64_t> batches_{0};
    std::thread worker_;
};

}  // namespace cluster
