// cluster_test.cpp — the reference's own tests for the dispatch path, re-stated against the
// C++ mirror (ptype_b200/host/cluster.hpp).  Each test names the Go test it follows.
//   ./cluster_test cpu   registry tests (no GPU)
//   ./cluster_test gpu   + client / balancer tests through the C ABI on cuda:0
#include <cstdio>
#include <cstdlib>
#include <set>

#include "../../ptype_b200/host/cluster.hpp"

using namespace cluster;
using namespace std::chrono_literals;

static int g_fail = 0, g_run = 0;
#define REQUIRE(cond)                                                    \
    do {                                                                 \
        if (!(cond)) {                                                   \
            std::printf("  FAIL %s:%d: %s\n", __FILE__, __LINE__, #cond); \
            ++g_fail;                                                    \
            return;                                                      \
        }                                                                \
    } while (0)
#define RUN(fn)                       \
    do {                              \
        std::printf("%s\n", #fn);     \
        ++g_run;                      \
        fn();                         \
    } while (0)

// mockRegistry (cluster/rpc_test.go:16-40): WatchService hands out a channel the test feeds.
struct MockRegistry : Registry {
    std::shared_ptr<NodeChan> nodeChan = std::make_shared<NodeChan>();
    bool sendInitial = false;
    std::vector<Node> initial;
    explicit MockRegistry(std::vector<Node> init) : sendInitial(true), initial(std::move(init)) {}
    MockRegistry() = default;
    Error Register(const std::string&, const std::string&, const std::string&, int) override { return Error(); }
    Error Services(std::map<std::string, std::vector<Node>>* out) override {
        out->clear();
        return Error();
    }
    std::shared_ptr<NodeChan> WatchService(const std::string&) override {
        if (sendInitial) nodeChan->Send(initial);
        return nodeChan;
    }
    void CancelWatch(const std::shared_ptr<NodeChan>& ch) override { ch->Close(); }
};

static ConnConfig testConnConfig() {  // rpc_test.go:79-83 (note Retries == 0 there)
    ConnConfig c;
    c.MaxConnections = DefaultConnConfig().MaxConnections;
    c.InitialNodeTimeout = 300ms;
    c.DebounceTime = 150ms;
    c.Retries = 0;
    return c;
}

// ------------------------------------------------------------------ registry (CPU)

// TestEtcdRegistry_Register (registry_test.go:41-87)
static void TestRegistry_Register() {
    MemRegistry sr;
    REQUIRE(!sr.Register("foo", "node1", "host", 8000));
    REQUIRE(!sr.Register("foo", "node2", "host2", 8000));
    REQUIRE(!sr.Register("bar", "node3", "host3", 3000));
    std::vector<Node> n;
    REQUIRE(!sr.nodes("foo", &n));
    REQUIRE((n == std::vector<Node>{{"host", 8000}, {"host2", 8000}}));
    REQUIRE(!sr.nodes("bar", &n));
    REQUIRE((n == std::vector<Node>{{"host3", 3000}}));
    REQUIRE(detail::MarshalNode(Node{"host", 8000}) == "{\"address\":\"host\",\"port\":8000}");
}

// TestEtcdRegistry_Services (registry_test.go:90-122)
static void TestRegistry_Services() {
    MemRegistry sr;
    sr.Put("services/foo/node1/", "{\"address\":\"host\", \"port\":8000}");
    sr.Put("services/foo/node2/", "{\"address\":\"host2\", \"port\":8000}");
    sr.Put("services/bar/node3/", "{\"address\":\"host3\", \"port\":3000}");
    std::map<std::string, std::vector<Node>> got;
    REQUIRE(!sr.Services(&got));
    REQUIRE(got.size() == 2);
    REQUIRE((got["foo"] == std::vector<Node>{{"host", 8000}, {"host2", 8000}}));
    REQUIRE((got["bar"] == std::vector<Node>{{"host3", 3000}}));
}

// TestEtcdRegistry_WatchService (registry_test.go:149-192)
static void TestRegistry_WatchService() {
    MemRegistry sr;
    sr.Put("services/foo/node1/", "{\"address\":\"host\", \"port\":8000}");
    auto ch = sr.WatchService("foo");
    bool ok;
    auto v = ch->Recv(1s, &ok);
    REQUIRE(v && (*v == std::vector<Node>{{"host", 8000}}));
    sr.Put("services/foo/node3/", "{\"address\":\"host3\", \"port\":3000}");
    v = ch->Recv(1s, &ok);
    REQUIRE(v && (*v == std::vector<Node>{{"host", 8000}, {"host3", 3000}}));
    REQUIRE(sr.Delete("services/foo/node1/") == 1);
    v = ch->Recv(1s, &ok);
    REQUIRE(v && (*v == std::vector<Node>{{"host3", 3000}}));
    // ..._stops_with_context_cancel (registry_test.go:194-207)
    sr.CancelWatch(ch);
    v = ch->Recv(1s, &ok);
    REQUIRE(!v && !ok);
}

// TestEtcdRegistry_nodes (registry_test.go:209-236)
static void TestRegistry_nodes_prefix_case() {
    MemRegistry sr;
    sr.Put("services/foo/node1/", "{\"address\":\"host\", \"port\":8000}");
    sr.Put("services/foo/node2/", "{\"address\":\"host2\", \"port\":8000}");
    sr.Put("services/foo_broken_prefix_case/node3/", "{\"address\":\"host3\", \"port\":3000}");
    std::vector<Node> n;
    REQUIRE(!sr.nodes("foo", &n));
    REQUIRE((n == std::vector<Node>{{"host", 8000}, {"host2", 8000}}));
}

// TestKVGet / GetErrorsOnNoKey / GetWithPrefix / GetWithMultipleOptions / Put / Delete /
// DeleteNoKey (cluster/store_test.go:17-169)
static void TestKVStore() {
    KVStore kvs;
    std::vector<std::string> vals;
    // store_test.go: keys live under store/, a prefix Get returns values in key order
    REQUIRE(!kvs.Put("raccoon1/", "uwu1") && !kvs.Put("raccoon2/", "uwu2"));
    REQUIRE(!kvs.Get("raccoon1", &vals, true));
    REQUIRE((vals == std::vector<std::string>{"uwu1"}));
    REQUIRE(kvs.Get("raccoon", &vals).Is(ErrNoKey()) && vals.empty());
    REQUIRE(!kvs.Get("raccoon", &vals, true));
    REQUIRE((vals == std::vector<std::string>{"uwu1", "uwu2"}));
    REQUIRE(!kvs.Put("hello", "world"));
    REQUIRE(!kvs.Get("hello", &vals) && vals == std::vector<std::string>{"world"});
    // registry images by epoch; the latest epoch wins
    std::string blob;
    REQUIRE(kvs.LatestImage(&blob).Is(ErrNoKey()));
    REQUIRE(!kvs.PublishImage(9, "nine", 4) && !kvs.PublishImage(10, "ten!", 4) && !kvs.PublishImage(2, "two.", 4));
    REQUIRE(!kvs.LatestImage(&blob) && blob == "ten!");
}

// example/optimus/coordinator/coordinator.go:67-99
static void TestOptimus_splitWork_watchReplies() {
    auto w = optimus::SplitWork(91);
    REQUIRE(w.size() == 10 && w[0].Min == 2 && w[0].Max == 10 && w[1].Min == 10 && w[9].Max == 100);
    REQUIRE(optimus::SplitWork(0).empty() && optimus::SplitWork(-7).empty() && optimus::SplitWork(10).size() == 1);
    REQUIRE(optimus::WatchReplies(91, {91, 91, 7, 13}) == 7);
    REQUIRE(optimus::WatchReplies(97, {97, 97, 97}) == 97);
}

// ------------------------------------------------------------- client / balancer (GPU)

static std::vector<Node> fakeNodes(int n, int base = 40000) {
    std::vector<Node> v;
    for (int i = 0; i < n; ++i) v.push_back(Node{"127.0.0.1", base + i});
    return v;
}

// TestClient_default_conn_config (rpc_test.go:85-90): an empty initial list is accepted,
// and a later Call yields ErrNoClientAvailable (rpc.go:62-64,178-180)
static void TestClient_default_conn_config() {
    Dispatcher gpu("", DefaultConnConfig().MaxConnections);
    REQUIRE(!gpu.Err());
    MockRegistry mock{std::vector<Node>{}};
    std::unique_ptr<Client> c;
    ConnConfig fast = DefaultConnConfig();
    fast.Retries = 0;
    REQUIRE(!Client::New(&gpu, "foo", &mock, &fast, &c));
    int64_t reply = 123;
    const int64_t args[3] = {7, 8, 0};
    Error e = c->Call("Calculator.Multiply", args, &reply);
    REQUIRE(e.Is(ErrNoClientAvailable()));
    REQUIRE(reply == 123);
}

// TestClient_Call (rpc_test.go:92-116) with the calculator example's method (client.go:38-44)
static void TestClient_Call() {
    Dispatcher gpu("", 3);
    gpu.Expose("foo", "Calculator");
    MockRegistry mock{fakeNodes(1)};
    ConnConfig cfg = testConnConfig();
    std::unique_ptr<Client> c;
    REQUIRE(!Client::New(&gpu, "foo", &mock, &cfg, &c));
    int64_t reply = 0;
    const int64_t args[3] = {7, 8, 0};
    REQUIRE(!c->Call("Calculator.Multiply", args, &reply));
    REQUIRE(reply == 56);
    Error e = c->Call("Calculator.Divide", args, &reply);
    REQUIRE(e.msg == "rpc: can't find method Calculator.Divide" && reply == 56);
    e = c->Call("Prime.Check", args, &reply);
    REQUIRE(e.msg == "rpc: can't find service Prime.Check");
    e = c->Call("nodot", args, &reply);
    REQUIRE(e.msg == "rpc: service/method request ill-formed: nodot");
}

// TestClient_Call_with_retry_multi_node (rpc_test.go:118-158): with 2 nodes the FIRST attempt
// goes to index 1 (AddUint64 returns the new value)
static void TestClient_first_call_goes_to_index_1() {
    Dispatcher gpu("", 3);
    gpu.Expose("foo", "Calculator");
    MockRegistry mock{fakeNodes(2)};
    ConnConfig cfg = testConnConfig();
    std::unique_ptr<Client> c;
    REQUIRE(!Client::New(&gpu, "foo", &mock, &cfg, &c));
    BatchReply rep[4];
    const int64_t args[4][3] = {{1, 1, 0}, {2, 2, 0}, {3, 3, 0}, {4, 4, 0}};
    REQUIRE(!c->CallBatch("Calculator.Add", args, 4, rep));
    auto sel = c->getSelectedNodes();
    REQUIRE(sel.size() == 2);
    REQUIRE(gpu.NodeOf(rep[0].NodeId) == sel[1]);
    REQUIRE(gpu.NodeOf(rep[1].NodeId) == sel[0]);
    REQUIRE(gpu.NodeOf(rep[2].NodeId) == sel[1]);
    REQUIRE(rep[3].Value == 8);
}

// TestClient_Call_with_retry_error (rpc_test.go:160-186): one node, the method fails, Retries
// is 0 in testConnConfig -> one attempt, an error, and the caller's reply is still 0
static void TestClient_Call_with_retry_error() {
    Dispatcher gpu("", 3);
    gpu.Expose("foo", "Prime");
    MockRegistry mock{fakeNodes(1)};
    ConnConfig cfg = testConnConfig();
    REQUIRE(cfg.Retries == 0);
    std::unique_ptr<Client> c;
    REQUIRE(!Client::New(&gpu, "foo", &mock, &cfg, &c));
    int64_t calls = 0;
    const int64_t wide[3] = {2, int64_t(1) << 40, int64_t(1) << 41};   // a Check the device refuses to walk
    Error e = c->Call("Prime.Check", wide, &calls);
    REQUIRE(e && calls == 0);
    REQUIRE(c->Seq() == 1);   // the one attempt drew one sequence number
}

// withRetry (rpc.go:107-116) + roundRobinSelect: every attempt draws the next sequence number,
// so after a call that failed 1 + Retries times the next call is routed by seq 1 + Retries + 1.
static void TestClient_retry_draws_a_sequence_number_per_attempt() {
    Dispatcher gpu("", 3);
    gpu.Expose("foo", "Calculator");
    MockRegistry mock{fakeNodes(3)};
    ConnConfig cfg = testConnConfig();
    cfg.Retries = 2;
    std::unique_ptr<Client> c;
    REQUIRE(!Client::New(&gpu, "foo", &mock, &cfg, &c));
    int64_t reply = 0;
    const int64_t args[3] = {5, 6, 0};
    REQUIRE(c->Call("Calculator.Nope", args, &reply).msg == "rpc: can't find method Calculator.Nope");
    REQUIRE(reply == 0 && c->Seq() == 3);
    BatchReply rep[1];
    const int64_t one[1][3] = {{5, 6, 0}};
    REQUIRE(!c->CallBatch("Calculator.Multiply", one, 1, rep));
    auto sel = c->getSelectedNodes();
    REQUIRE(sel.size() == 3 && rep[0].Value == 30 && c->Seq() == 4);
    REQUIRE(gpu.NodeOf(rep[0].NodeId) == sel[4 % 3]);
}

// TestNewConnectionBalancer_successul_initial_connect (rpc_test.go:293-304)
static void TestBalancer_successful_initial_connect() {
    Dispatcher gpu("", 3);
    auto n = fakeNodes(1);
    MockRegistry mock{n};
    ConnConfig cfg = testConnConfig();
    std::unique_ptr<Client> c;
    REQUIRE(!Client::New(&gpu, "foo", &mock, &cfg, &c));
    REQUIRE((c->getSelectedNodes() == std::vector<Node>{n[0]}));
    REQUIRE(!c->ConnectionErrs().has_value());
}

// TestClient_Go / TestClient_Go_with_retry_error (rpc_test.go:188-215,262-292): async call,
// completion signalled; on failure the call carries the error and the reply stays 0
static void TestClient_Go() {
    Dispatcher gpu("", 3);
    gpu.Expose("foo", "Calculator");
    MockRegistry mock{fakeNodes(1)};
    ConnConfig cfg = testConnConfig();
    std::unique_ptr<Client> c;
    REQUIRE(!Client::New(&gpu, "foo", &mock, &cfg, &c));
    int64_t reply = 0, untouched = 0;
    const int64_t args[3] = {7, 8, 0};
    auto call = c->Go("Calculator.Multiply", args, &reply);
    auto bad = c->Go("Calculator.Nope", args, &untouched);
    call->Wait();
    bad->Wait();
    REQUIRE(call->Done() && !call->Err && reply == 56);
    REQUIRE(bad->Err.msg == "rpc: can't find method Calculator.Nope" && untouched == 0);
    // the same through the dispatch loop: many asynchronous calls share GPU batches, none retried
    cfg.Retries = 2;
    std::unique_ptr<Client> c2;
    REQUIRE(!Client::New(&gpu, "foo", &mock, &cfg, &c2));
    CallCoalescer loop(c2.get());
    std::vector<int64_t> replies(500, -1);
    std::vector<std::shared_ptr<Client::AsyncCall>> calls;
    for (int i = 0; i < 500; ++i) {
        const int64_t a[3] = {i, 3, 0};
        calls.push_back(loop.Go(i % 50 == 49 ? "Calculator.Nope" : "Calculator.Multiply", a, &replies[i]));
    }
    for (auto& k : calls) k->Wait();
    for (int i = 0; i < 500; ++i) {
        if (i % 50 == 49)
            REQUIRE(calls[i]->Err.msg == "rpc: can't find method Calculator.Nope" && replies[i] == -1);
        else
            REQUIRE(!calls[i]->Err && replies[i] == 3 * i);
    }
    REQUIRE(c2->Seq() == 500);          // one attempt per call although Retries is 2 (rpc.go:91)
    REQUIRE(loop.Batches() < 100);      // gathered, not one launch per call
}

// The optimus example end to end: one target's splitWork stream through a Prime client, the
// answer picked by watchReplies (coordinator.go:46-99; the 250 ms sleep of prime.go:17 dropped)
static void TestExample_optimus() {
    MemRegistry reg;
    REQUIRE(!reg.Register("prime_worker", "prime_worker_1", "127.0.0.1", 8000));
    REQUIRE(!reg.Register("prime_worker", "prime_worker_2", "127.0.0.1", 8002));
    Dispatcher gpu("127.0.0.1", 3);
    gpu.Expose("prime_worker", "Prime");
    ConnConfig cfg = testConnConfig();
    std::unique_ptr<Client> c;
    REQUIRE(!Client::New(&gpu, "prime_worker", &reg, &cfg, &c));
    for (int64_t target : {91, 97, 100, 121, 7919, 7917}) {
        auto work = optimus::SplitWork(target);
        std::vector<std::array<int64_t, 3>> args;
        for (auto& a : work) args.push_back({a.Min, a.Max, a.Target});
        std::vector<BatchReply> rep(args.size());
        REQUIRE(!c->CallBatch("Prime.Check", reinterpret_cast<const int64_t(*)[3]>(args.data()), args.size(), rep.data()));
        std::vector<int64_t> values;
        for (auto& r : rep) {
            REQUIRE(r.Status == 0);
            values.push_back(r.Value);
        }
        const int64_t answer = optimus::WatchReplies(target, values);
        int64_t smallest = target;
        for (int64_t d = 2; d < target; ++d)
            if (target % d == 0) {
                smallest = d;
                break;
            }
        REQUIRE(answer == smallest);   // in order of submission the first non-target reply is the smallest divisor
    }
}

// Client.Call is goroutine-safe (rpc.go:59,181,287-291): many concurrent callers, each gets
// its own reply; here they are gathered into batches by the CallCoalescer.  The sequence
// numbers handed out are exactly 1..N (one counter), so the 3 selected nodes are used evenly.
static void TestCoalescer_concurrent_callers() {
    Dispatcher gpu("10.0.0.5", 3);
    gpu.Expose("calculator", "Calculator");
    MemRegistry reg;
    for (int i = 1; i <= 16; ++i) {
        char name[32], addr[32];
        std::snprintf(name, sizeof name, "calculator_node_%02d", i);
        std::snprintf(addr, sizeof addr, "10.0.0.%d", i);
        REQUIRE(!reg.Register("calculator", name, addr, 31234));
    }
    ConnConfig cfg = testConnConfig();
    std::unique_ptr<Client> c;
    REQUIRE(!Client::New(&gpu, "calculator", &reg, &cfg, &c));
    CallCoalescer co(c.get(), 4096, std::chrono::microseconds(300));
    constexpr int kThreads = 16, kCalls = 1500;
    std::atomic<int> wrong{0}, errors{0};
    std::vector<std::thread> ts;
    for (int t = 0; t < kThreads; ++t)
        ts.emplace_back([&, t] {
            for (int i = 0; i < kCalls; ++i) {
                const int64_t a = t * 100000 + i, b = i - 700;
                const int64_t args[3] = {a, b, 0};
                int64_t reply = -1;
                const bool mul = (i + t) % 2 == 0;
                Error e = co.Call(mul ? "Calculator.Multiply" : "Calculator.Add", args, &reply);
                if (e) ++errors;
                if (reply != (mul ? a * b : a + b)) ++wrong;
                if (i % 500 == 0) {
                    int64_t untouched = 7;
                    if (co.Call("Calculator.Nope", args, &untouched).msg != "rpc: can't find method Calculator.Nope" ||
                        untouched != 7)
                        ++wrong;
                }
            }
        });
    for (auto& th : ts) th.join();
    REQUIRE(errors == 0 && wrong == 0);
    REQUIRE(co.Batches() < static_cast<uint64_t>(kThreads) * kCalls / 2);   // calls really were coalesced
}

// A batch that mixes services: every service's Client draws from its own counter, in batch
// order, and a service without nodes draws nothing (rpc.go:129,172-183) - on the device.
static void TestDispatcher_mixed_batch_draws_seq_on_device() {
    Dispatcher gpu("", 3);
    gpu.Expose("foo", "Calculator");
    gpu.Expose("bar", "Calculator");
    auto n = fakeNodes(5);
    REQUIRE(!gpu.SetNodes("foo", {n[0], n[1]}));
    REQUIRE(!gpu.SetNodes("bar", {n[2], n[3], n[4]}));
    REQUIRE(!gpu.SetNodes("idle", {}));
    uint32_t method = 0, recv = 0;
    ptype_b200_method_key("Calculator.Multiply", &method, &recv);
    const char* names[3] = {"foo", "bar", "idle"};
    std::vector<ptype_req> reqs(3000);
    std::vector<int> which(reqs.size());
    for (size_t j = 0; j < reqs.size(); ++j) {
        which[j] = static_cast<int>((j * 7 + j / 5) % 3);
        reqs[j] = ptype_req{};
        reqs[j].svc_key = gpu.ServiceKey(names[which[j]]);
        reqs[j].seq = 999;  // ignored
        reqs[j].method = method;
        reqs[j].recv = recv;
        reqs[j].a0 = static_cast<int64_t>(j);
        reqs[j].a1 = 3;
    }
    std::vector<ptype_reply> rep;
    REQUIRE(!gpu.CallMixed(reqs, &rep));
    const std::vector<Node> sel[2] = {gpu.SelectedNodes("foo"), gpu.SelectedNodes("bar")};
    REQUIRE(sel[0].size() == 2 && sel[1].size() == 3);
    uint64_t drawn[3] = {0, 0, 0};
    int bad = 0;
    for (size_t j = 0; j < reqs.size(); ++j) {
        const int s = which[j];
        if (s == 2) {   // no nodes: ErrNoClientAvailable, nothing drawn
            if (rep[j].status != PTYPE_ST_NO_CLIENT || reqs[j].seq != 0) ++bad;
            continue;
        }
        const uint64_t seq = ++drawn[s];
        if (reqs[j].seq != seq || rep[j].status != PTYPE_ST_OK || rep[j].value != static_cast<int64_t>(j) * 3 ||
            !(gpu.NodeOf(rep[j].node) == sel[s][seq % sel[s].size()]))
            ++bad;
    }
    REQUIRE(bad == 0);
    REQUIRE(gpu.Seq("foo") == drawn[0] && gpu.Seq("bar") == drawn[1] && gpu.Seq("idle") == 0);
    // a node joins foo: new image, new selection, the counters carry on
    REQUIRE(!gpu.SetNodes("foo", {n[0], n[1], n[4]}));
    std::vector<ptype_req> more(10, reqs[0]);
    for (auto& q : more) q.svc_key = gpu.ServiceKey("foo");
    REQUIRE(!gpu.CallMixed(more, &rep));
    REQUIRE(more[0].seq == drawn[0] + 1 && more[9].seq == drawn[0] + 10);
}

// TestNewConnectionBalancer_with_no_available_server (rpc_test.go:307-314)
static void TestNewClient_with_no_available_server() {
    Dispatcher gpu("", 3);
    MockRegistry mock;  // never sends
    ConnConfig cfg = testConnConfig();
    std::unique_ptr<Client> c;
    Error e = Client::New(&gpu, "foo", &mock, &cfg, &c);
    REQUIRE(e.msg == "no initial nodes provided for foo");
    REQUIRE(!c);
}

// TestConnectionBalancer_watchForNewNodes (rpc_test.go:323-388)
static void TestBalancer_watchForNewNodes() {
    Dispatcher gpu("", 3);
    auto n = fakeNodes(4);
    Node node = n[0], node2 = n[1], node3 = n[2], node4 = n[3];
    MockRegistry mock{std::vector<Node>{node}};
    ConnConfig cfg = testConnConfig();
    std::unique_ptr<Client> c;
    REQUIRE(!Client::New(&gpu, "foo", &mock, &cfg, &c));
    REQUIRE((c->getSelectedNodes() == std::vector<Node>{node}));
    // more than max connections of nodes get sent -> [node4, node, node2]
    mock.nodeChan->Send({node, node2, node3, node4});
    REQUIRE(c->WaitConnsUpdated(2, 3s));
    REQUIRE((c->getSelectedNodes() == std::vector<Node>{node4, node, node2}));
    // connected node is removed -> identity
    mock.nodeChan->Send({node, node3, node4});
    REQUIRE(c->WaitConnsUpdated(3, 3s));
    REQUIRE((c->getSelectedNodes() == std::vector<Node>{node, node3, node4}));
    // debounce: only the last of a burst is applied, once; empty lists are ignored
    mock.nodeChan->Send({node});
    mock.nodeChan->Send({node2});
    mock.nodeChan->Send({});
    mock.nodeChan->Send({node3});
    mock.nodeChan->Send({node, node2, node3});
    REQUIRE(c->WaitConnsUpdated(4, 3s));
    REQUIRE((c->getSelectedNodes() == std::vector<Node>{node, node2, node3}));
    std::this_thread::sleep_for(400ms);
    REQUIRE(c->ConnsUpdated() == 4);
    REQUIRE(!c->ConnectionErrs());
}

// TestConnectionBalancer_roundRobinSelect (rpc_test.go:390-425): n consecutive selections are
// n distinct clients, and the sequence wraps
static void TestBalancer_roundRobinSelect() {
    Dispatcher gpu("", 0);  // mesh: all 100 nodes selected
    gpu.Expose("foo", "Calculator");
    MockRegistry mock{fakeNodes(100)};
    ConnConfig cfg = testConnConfig();
    cfg.MaxConnections = 0;
    std::unique_ptr<Client> c;
    REQUIRE(!Client::New(&gpu, "foo", &mock, &cfg, &c));
    std::vector<BatchReply> rep(200);
    std::vector<std::array<int64_t, 3>> args(200, std::array<int64_t, 3>{1, 2, 0});
    REQUIRE(!c->CallBatch("Calculator.Add", reinterpret_cast<const int64_t(*)[3]>(args.data()), 200, rep.data()));
    std::set<uint32_t> first, second;
    for (int i = 0; i < 100; ++i) first.insert(rep[i].NodeId);
    for (int i = 100; i < 200; ++i) second.insert(rep[i].NodeId);
    REQUIRE(first.size() == 100 && second.size() == 100);
    for (int i = 0; i < 100; ++i) REQUIRE(rep[i].NodeId == rep[i + 100].NodeId);
    for (auto& r : rep) REQUIRE(r.Status == 0 && r.Value == 3);
}

// TestConnectionBalancer_mesh_network (rpc_test.go:427-476)
static void TestBalancer_mesh_network() {
    Dispatcher gpu("", 0);
    auto n = fakeNodes(4);
    MockRegistry mock{std::vector<Node>{n[0], n[1]}};
    ConnConfig cfg = testConnConfig();
    cfg.MaxConnections = 0;
    std::unique_ptr<Client> c;
    REQUIRE(!Client::New(&gpu, "foo", &mock, &cfg, &c));
    REQUIRE((c->getSelectedNodes() == std::vector<Node>{n[0], n[1]}));
    mock.nodeChan->Send({n[0], n[1], n[2]});
    REQUIRE(c->WaitConnsUpdated(2, 3s));
    REQUIRE((c->getSelectedNodes() == std::vector<Node>{n[0], n[1], n[2]}));
    mock.nodeChan->Send({n[0], n[1], n[2], n[3]});
    REQUIRE(c->WaitConnsUpdated(3, 3s));
    REQUIRE((c->getSelectedNodes() == n));
}

// Registry -> Client end to end with the in-memory registry: the calculator example
// (server.go:16-38 registers, client.go:33-44 calls), then a node leaves and calls re-route.
static void TestExample_calculator_with_registry() {
    MemRegistry reg;
    for (int i = 1; i <= 16; ++i) {
        char name[32], addr[32];
        std::snprintf(name, sizeof name, "calculator_node_%02d", i);
        std::snprintf(addr, sizeof addr, "10.0.0.%d", i);
        REQUIRE(!reg.Register("calculator", name, addr, 31234));
    }
    Dispatcher gpu("10.0.0.5", 3);
    gpu.Expose("calculator", "Calculator");
    ConnConfig cfg = testConnConfig();
    cfg.Retries = 2;
    std::unique_ptr<Client> c;
    REQUIRE(!Client::New(&gpu, "calculator", &reg, &cfg, &c));
    // derived known answer: localAddr 10.0.0.5, 16 nodes, max 3 -> [3,0,9]
    auto sel = c->getSelectedNodes();
    REQUIRE((sel == std::vector<Node>{{"10.0.0.4", 31234}, {"10.0.0.1", 31234}, {"10.0.0.10", 31234}}));
    int64_t reply = 0;
    const int64_t args[3] = {7, 8, 0};
    REQUIRE(!c->Call("Calculator.Multiply", args, &reply));
    REQUIRE(reply == 56);
    const int64_t wrap[3] = {INT64_C(3037000500), INT64_C(3037000500), 0};
    REQUIRE(!c->Call("Calculator.Multiply", wrap, &reply));
    REQUIRE(reply == INT64_C(-9223372036709301616));
    REQUIRE(reg.Delete("services/calculator/calculator_node_04/") == 1);
    REQUIRE(c->WaitConnsUpdated(2, 3s));
    sel = c->getSelectedNodes();
    REQUIRE(sel.size() == 3 && sel[0] != Node({"10.0.0.4", 31234}));
}

// The batch call on reusable page-locked buffers (what the bench's e2e_mirror times): a mixed
// Add/Multiply batch routed like consecutive Call()s, values in place, failures listed, and
// withRetry re-issuing only the calls that failed - each under a fresh sequence number.
static void TestClient_CallBatch_on_pinned_buffers() {
    Dispatcher gpu("", 3);
    gpu.Expose("foo", "Calculator");
    MockRegistry mock{fakeNodes(5)};
    ConnConfig cfg = testConnConfig();
    cfg.Retries = 2;
    std::unique_ptr<Client> c;
    REQUIRE(!Client::New(&gpu, "foo", &mock, &cfg, &c));
    const size_t n = 100003;
    auto b = c->NewBatch(n, 2);
    REQUIRE(b->Ok() && b->Capacity() == n && b->ArgsPerCall() == 2);
    for (size_t j = 0; j < n; ++j) {
        b->Args()[2 * j] = static_cast<int64_t>(j) - 50000;
        b->Args()[2 * j + 1] = INT64_C(0x7FFFFFFFFFFF) * static_cast<int64_t>(j % 7);
        b->Sel()[j] = static_cast<uint8_t>(j % 3 == 0);   // 1 = Multiply
    }
    REQUIRE(!c->CallBatch({"Calculator.Add", "Calculator.Multiply"}, *b, n));
    REQUIRE(b->Failed().empty() && c->Seq() == n);
    for (size_t j = 0; j < n; ++j) {
        const uint64_t x = static_cast<uint64_t>(b->Args()[2 * j]), y = static_cast<uint64_t>(b->Args()[2 * j + 1]);
        REQUIRE(b->Values()[j] == static_cast<int64_t>(b->Sel()[j] ? x * y : x + y));
    }
    // one method of the three is unknown: those calls fail on every attempt (1 + Retries of them),
    // the others succeed on the first, and every attempt of every call drew its own number
    for (size_t j = 0; j < n; ++j) b->Sel()[j] = static_cast<uint8_t>(j % 3);
    REQUIRE(!c->CallBatch({"Calculator.Add", "Calculator.Multiply", "Calculator.Divide"}, *b, n));
    const size_t bad = n / 3;  // j % 3 == 2
    REQUIRE(b->Failed().size() == bad && c->Seq() == n + n + 2 * bad);
    for (const ptype_failed& f : b->Failed())
        REQUIRE(f.index % 3 == 2 && f.status == PTYPE_ST_CANT_FIND_METHOD && b->Values()[f.index] == 0);
    // a batch smaller than the buffers, reused; an empty one; one that does not fit
    REQUIRE(!c->CallBatch({"Calculator.Add"}, *b, 10));
    REQUIRE(b->Failed().empty() && b->Values()[9] == b->Args()[18] + b->Args()[19]);
    REQUIRE(!c->CallBatch({"Calculator.Add"}, *b, 0));
    REQUIRE(c->CallBatch({"Calculator.Add"}, *b, n + 1));
    // the array form still reports the routing decision of every call
    BatchReply rep[3];
    const int64_t args[3][3] = {{1, 2, 0}, {3, 4, 0}, {5, 6, 0}};
    const uint64_t before = c->Seq();
    REQUIRE(!c->CallBatch("Calculator.Multiply", args, 3, rep));
    auto sel = c->getSelectedNodes();
    for (uint64_t t = 0; t < 3; ++t)
        REQUIRE(rep[t].Status == 0 && gpu.NodeOf(rep[t].NodeId) == sel[(before + 1 + t) % 3]);
    REQUIRE(rep[2].Value == 30);
}

// one host builds the snapshot and publishes it through the KV store, another
// host (its own Dispatcher and device context) installs the newest epoch and answers alike.
static void TestDispatcher_image_through_the_store() {
    KVStore store;
    Dispatcher builder("10.0.0.5", 3), follower("10.0.0.5", 3);
    REQUIRE(follower.LoadLatestImage(&store).Is(ErrNoKey()));
    builder.Expose("calculator", "Calculator");
    MemRegistry reg;
    for (int i = 1; i <= 16; ++i) {
        char name[32], addr[32];
        std::snprintf(name, sizeof name, "calculator_node_%02d", i);
        std::snprintf(addr, sizeof addr, "10.0.0.%d", i);
        REQUIRE(!reg.Register("calculator", name, addr, 31234));
    }
    ConnConfig cfg = testConnConfig();
    std::unique_ptr<Client> c;
    REQUIRE(!Client::New(&builder, "calculator", &reg, &cfg, &c));
    REQUIRE(!builder.PublishImage(&store));
    REQUIRE(!follower.LoadLatestImage(&store));
    REQUIRE(follower.Epoch() == builder.Epoch());
    REQUIRE(follower.SelectedNodes("calculator") == builder.SelectedNodes("calculator"));
    // the same 48-byte records through either context: the same replies
    std::vector<ptype_req> reqs(1000);
    uint32_t m = 0, r = 0;
    ptype_b200_method_key("Calculator.Multiply", &m, &r);
    for (size_t j = 0; j < reqs.size(); ++j)
        reqs[j] = ptype_req{builder.ServiceKey("calculator"), j + 1, m, r, static_cast<int64_t>(j), 3, 0};
    std::vector<ptype_reply> a(reqs.size()), b(reqs.size());
    REQUIRE(ptype_b200_dispatch(builder.ctx(), 0, reqs.data(), reqs.size(), a.data(), 0) == PTYPE_OK);
    REQUIRE(ptype_b200_dispatch(follower.ctx(), 0, reqs.data(), reqs.size(), b.data(), 0) == PTYPE_OK);
    REQUIRE(std::memcmp(a.data(), b.data(), a.size() * sizeof(ptype_reply)) == 0);
    REQUIRE(a[7].value == 21 && a[7].status == PTYPE_ST_OK);
    // a node leaves: new epoch on the builder, published, picked up by the follower
    REQUIRE(reg.Delete("services/calculator/calculator_node_04/") == 1);
    REQUIRE(c->WaitConnsUpdated(2, 3s));
    REQUIRE(!builder.PublishImage(&store));
    REQUIRE(!follower.LoadLatestImage(&store));
    REQUIRE(follower.Epoch() == builder.Epoch() && follower.Epoch() == 2);
    REQUIRE(follower.SelectedNodes("calculator") == builder.SelectedNodes("calculator"));
}

int main(int argc, char** argv) {
    const bool gpu = argc > 1 && std::string(argv[1]) == "gpu";
    RUN(TestRegistry_Register);
    RUN(TestRegistry_Services);
    RUN(TestRegistry_WatchService);
    RUN(TestRegistry_nodes_prefix_case);
    RUN(TestKVStore);
    RUN(TestOptimus_splitWork_watchReplies);
    if (gpu) {
        RUN(TestClient_Go);
        RUN(TestCoalescer_concurrent_callers);
        RUN(TestDispatcher_mixed_batch_draws_seq_on_device);
        RUN(TestExample_optimus);
        RUN(TestClient_default_conn_config);
        RUN(TestClient_Call);
        RUN(TestClient_first_call_goes_to_index_1);
        RUN(TestClient_CallBatch_on_pinned_buffers);
        RUN(TestClient_Call_with_retry_error);
        RUN(TestClient_retry_draws_a_sequence_number_per_attempt);
        RUN(TestBalancer_successful_initial_connect);
        RUN(TestNewClient_with_no_available_server);
        RUN(TestBalancer_watchForNewNodes);
        RUN(TestBalancer_roundRobinSelect);
        RUN(TestBalancer_mesh_network);
        RUN(TestExample_calculator_with_registry);
        RUN(TestDispatcher_image_through_the_store);
    }
    std::printf("%d tests, %d failed\n", g_run, g_fail);
    return g_fail ? 1 : 0;
}
