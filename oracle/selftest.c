/*
 * selftest.c — pins the oracle against the results the reference's own tests
 * fix for this path and against known answers derived
 * from the Go language rules.  TEST INFRASTRUCTURE ONLY.
 */
#include <inttypes.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "ptype_oracle.h"

static int failures = 0;
#define CHECK(cond)                                                        \
    do {                                                                   \
        if (!(cond)) {                                                     \
            printf("FAIL %s:%d: %s\n", __FILE__, __LINE__, #cond);         \
            failures++;                                                    \
        }                                                                  \
    } while (0)

static void expect_sel(const char* addr, size_t n, int64_t max, const uint32_t* want, long k) {
    uint32_t got[128];
    long g = po_select_nodes(addr, n, max, got, 128);
    CHECK(g == k);
    for (long i = 0; i < k && i < g; i++) CHECK(got[i] == want[i]);
}

int main(void) {
    /* hash/fnv known answers (re-derived in SURVEY.md §8c) */
    CHECK(po_fnv1a32("0", 1) == 0x350ca8afu);
    CHECK(po_fnv1a32("1", 1) == 0x340ca71cu);
    CHECK(po_fnv1a32("2", 1) == 0x370cabd5u);
    CHECK(po_fnv1a32("Calculator.Multiply", 19) == 0xa9293851u);
    CHECK(po_fnv1a32("Calculator.Add", 14) == 0x29baa940u);
    CHECK(po_fnv1a32("Prime.Check", 11) == 0x7238b75au);
    CHECK(po_fnv1a64_salted("calculator", 10, 0) == 0x2ea3a3998f4333adull);
    CHECK(po_fnv1a64_salted("prime_worker", 12, 0) == 0xe07ee02192b4a67full);

    /* G1 cluster/rpc_test.go:341-354: localAddr "", 4 nodes, max 3 → [n4,n1,n2] */
    {
        uint32_t w[] = {3, 0, 1};
        expect_sel("", 4, 3, w, 3);
    }
    /* G2 rpc_test.go:356-369: n <= max → identity */
    {
        uint32_t w[] = {0, 1, 2};
        expect_sel("", 3, 3, w, 3);
    }
    /* G4 rpc_test.go:427-476: mesh (max 0) → all nodes, for 2,3,4 nodes */
    for (size_t n = 2; n <= 4; n++) {
        uint32_t w[] = {0, 1, 2, 3};
        expect_sel("", n, 0, w, (long)n);
    }
    /* derived selections */
    {
        uint32_t a[] = {15, 12, 5}, b[] = {47, 28, 21}, c[] = {10, 13, 4}, d[] = {3, 0, 9},
                 e[] = {35, 16, 9}, f[] = {4, 3, 4};
        expect_sel("", 16, 3, a, 3);
        expect_sel("", 64, 3, b, 3);
        expect_sel("127.0.0.1", 16, 3, c, 3);
        expect_sel("10.0.0.5", 16, 3, d, 3);
        expect_sel("10.0.0.5", 64, 3, e, 3);
        expect_sel("10.0.0.231", 6, 3, f, 3); /* duplicate kept: no dedup (rpc.go:258-261) */
    }
    CHECK(po_select_nodes("", 4, -1, NULL, 0) == -1);

    /* G5 rpc_test.go:118-158: first attempt with 2 nodes goes to index 1 */
    CHECK(po_round_robin(1, 2) == 1);
    /* round robin n=3: 1,2,0,1,2,0 */
    {
        int64_t want[] = {1, 2, 0, 1, 2, 0};
        for (uint64_t s = 1; s <= 6; s++) CHECK(po_round_robin(s, 3) == want[s - 1]);
    }
    /* G6 rpc_test.go:409-424: n consecutive selections are n distinct clients */
    {
        int seen[100] = {0};
        for (uint64_t s = 101; s <= 200; s++) {
            int64_t i = po_round_robin(s, 100);
            CHECK(i >= 0 && !seen[i]);
            if (i >= 0) seen[i] = 1;
        }
    }
    /* G8 rpc.go:178-180: empty list → nil client */
    CHECK(po_round_robin(1, 0) == -1);
    /* int(index) goes negative at 2^63: panic unless the remainder is 0 */
    CHECK(po_round_robin(0x8000000000000000ull, 3) == -2);
    CHECK(po_round_robin(0x8000000000000000ull, 2) == 0);

    /* wrapping int64 (Go spec) */
    CHECK(po_multiply(7, 8) == 56); /* example/calculator/client/client.go:38-44 */
    CHECK(po_multiply((int64_t)1 << 62, 2) == INT64_MIN);
    CHECK(po_multiply(INT64_MAX, INT64_MAX) == 1);
    CHECK(po_multiply(INT64_MIN, -1) == INT64_MIN);
    CHECK(po_multiply(3037000500ll, 3037000500ll) == -9223372036709301616ll);
    CHECK(po_add(INT64_MAX, 1) == INT64_MIN);

    /* Prime.Check over the coordinator's chunks */
    {
        int64_t w91[] = {7, 13, 91, 91, 91, 91, 91, 91, 91, 91};
        int64_t w100[] = {2, 10, 20, 100, 100, 50, 100, 100, 100, 100};
        int64_t chunks[16][3];
        size_t c = po_split_work(91, chunks, 16);
        CHECK(c == 10);
        for (size_t i = 0; i < c; i++)
            CHECK(po_check(chunks[i][0], chunks[i][1], chunks[i][2]) == w91[i]);
        c = po_split_work(100, chunks, 16);
        CHECK(c == 10);
        CHECK(chunks[0][0] == 2 && chunks[0][1] == 10 && chunks[1][0] == 10 && chunks[9][1] == 100);
        for (size_t i = 0; i < c; i++)
            CHECK(po_check(chunks[i][0], chunks[i][1], chunks[i][2]) == w100[i]);
        c = po_split_work(97, chunks, 16);
        CHECK(c == 10);
        for (size_t i = 0; i < c; i++) CHECK(po_check(chunks[i][0], chunks[i][1], 97) == 97);
        c = po_split_work(121, chunks, 16);
        CHECK(c == 13);
        for (size_t i = 0; i < c; i++)
            CHECK(po_check(chunks[i][0], chunks[i][1], 121) == (i == 1 ? 11 : 121));
        CHECK(po_split_work(0, NULL, 0) == 0);
        CHECK(po_split_work(-5, NULL, 0) == 0);
        CHECK(po_split_work(1, NULL, 0) == 1);
        CHECK(po_split_work(10, NULL, 0) == 1);
        CHECK(po_split_work(11, NULL, 0) == 2);
    }
    CHECK(po_check(-5, 10, 20) == -5);
    CHECK(po_check(0, 3, 7) == 1);
    CHECK(po_check(-1, 5, INT64_MIN) == INT64_MIN); /* bound = min(5, MinInt64): empty loop */
    CHECK(po_check(-1, 5, 4) == -1);                /* 4 % -1 == 0 */

    /* G9 cluster/registry_test.go:90-122: Services() grouping and order */
    {
        po_kv* kv = po_kv_new();
        po_kv_put(kv, "services/foo/node1/", "{\"address\":\"host\", \"port\":8000}");
        po_kv_put(kv, "services/foo/node2/", "{\"address\":\"host2\", \"port\":8000}");
        po_kv_put(kv, "services/bar/node3/", "{\"address\":\"host3\", \"port\":3000}");
        po_service_list* sl;
        long ns = po_services(kv, &sl);
        CHECK(ns == 2);
        for (long i = 0; i < ns; i++) {
            if (!strcmp(sl[i].name, "foo")) {
                CHECK(sl[i].n == 2);
                CHECK(!strcmp(sl[i].nodes[0].address, "host") && sl[i].nodes[0].port == 8000);
                CHECK(!strcmp(sl[i].nodes[1].address, "host2") && sl[i].nodes[1].port == 8000);
            } else {
                CHECK(!strcmp(sl[i].name, "bar") && sl[i].n == 1);
                CHECK(!strcmp(sl[i].nodes[0].address, "host3") && sl[i].nodes[0].port == 3000);
            }
        }
        po_services_free(sl, (size_t)ns);
        /* registry_test.go:209-236: nodes("foo") excludes foo_broken_prefix_case */
        po_kv_put(kv, "services/foo_broken_prefix_case/node3/", "{\"address\":\"host3\", \"port\":3000}");
        po_node* nd;
        long nn = po_nodes(kv, "foo", &nd);
        CHECK(nn == 2);
        CHECK(!strcmp(nd[0].address, "host") && !strcmp(nd[1].address, "host2"));
        free(nd);
        po_kv_free(kv);
    }
    /* registry_test.go:41-87: Register writes the JSON shape under the etcd key */
    {
        po_kv* kv = po_kv_new();
        po_register(kv, "foo", "node1", "host", 8000);
        size_t first;
        CHECK(po_kv_range_prefix(kv, "services/foo/", &first) == 1);
        CHECK(!strcmp(po_kv_key_at(kv, first), "services/foo/node1/"));
        CHECK(!strcmp(po_kv_value_at(kv, first), "{\"address\":\"host\",\"port\":8000}"));
        po_kv_free(kv);
    }
    /* Client.Call end to end on strings */
    {
        po_kv* kv = po_kv_new();
        for (int i = 1; i <= 16; i++) {
            char nm[64], ad[64];
            snprintf(nm, sizeof nm, "calculator_node_%02d", i);
            snprintf(ad, sizeof ad, "10.0.0.%d", i);
            po_register(kv, "calculator", nm, ad, 31234);
        }
        const char* declared[] = {"ghost"};
        po_snapshot* s = po_snapshot_build(kv, "10.0.0.5", 3, declared, 1, 0);
        po_snapshot_expose(s, "calculator", "Calculator");
        uint32_t sel[8];
        CHECK(po_snapshot_selected(s, "calculator", sel, 8) == 3);
        CHECK(sel[0] == 3 && sel[1] == 0 && sel[2] == 9);
        ptype_reply r;
        int64_t a[3] = {7, 8, 0};
        po_call(s, "calculator", 1, "Calculator.Multiply", a, UINT64_MAX, &r);
        CHECK(r.status == PTYPE_ST_OK && r.value == 56 && r.node == 0);
        po_call(s, "calculator", 3, "Calculator.Add", a, UINT64_MAX, &r);
        CHECK(r.status == PTYPE_ST_OK && r.value == 15 && r.node == 3);
        po_call(s, "calculator", 2, "Calculator.Divide", a, UINT64_MAX, &r);
        CHECK(r.status == PTYPE_ST_CANT_FIND_METHOD && r.value == 0 && r.node == 9);
        po_call(s, "calculator", 2, "Prime.Check", a, UINT64_MAX, &r);
        CHECK(r.status == PTYPE_ST_CANT_FIND_SERVICE);
        po_call(s, "calculator", 2, "nodot", a, UINT64_MAX, &r);
        CHECK(r.status == PTYPE_ST_ILL_FORMED);
        po_call(s, "ghost", 1, "Calculator.Multiply", a, UINT64_MAX, &r);
        CHECK(r.status == PTYPE_ST_NO_CLIENT && r.node == PTYPE_NO_NODE);
        po_call(s, "nobody", 1, "Calculator.Multiply", a, UINT64_MAX, &r);
        CHECK(r.status == PTYPE_ST_NO_SERVICE);
        po_snapshot_free(s);
        po_kv_free(kv);
    }
    /* gob value encodings */
    {
        uint8_t b[16];
        CHECK(po_gob_int(7, b) == 1 && b[0] == 0x0e);
        CHECK(po_gob_int(8, b) == 1 && b[0] == 0x10);
        CHECK(po_gob_int(56, b) == 1 && b[0] == 0x70);
        CHECK(po_gob_int(-1, b) == 1 && b[0] == 0x01);
        CHECK(po_gob_int(64, b) == 2 && b[0] == 0xff && b[1] == 0x80);
        CHECK(po_gob_int(-65, b) == 2 && b[0] == 0xff && b[1] == 0x81);
        CHECK(po_gob_int(INT64_MAX, b) == 9 && b[0] == 0xf8 && b[8] == 0xfe);
        CHECK(po_gob_int(INT64_MIN, b) == 9 && b[0] == 0xf8 && b[8] == 0xff);
        CHECK(po_gob_uint(256, b) == 3 && b[0] == 0xfe && b[1] == 1 && b[2] == 0);
    }
    /* etcdKey / filepath.Join */
    {
        char k[128];
        const char* e1[] = {"services", "foo", "node1"};
        CHECK(po_etcd_key(k, sizeof k, e1, 3) > 0 && !strcmp(k, "services/foo/node1/"));
        const char* e2[] = {"services", "foo"};
        CHECK(po_etcd_key(k, sizeof k, e2, 2) > 0 && !strcmp(k, "services/foo/"));
        const char* e3[] = {"services", "", "a//b/../c"};
        CHECK(po_etcd_key(k, sizeof k, e3, 3) > 0 && !strcmp(k, "services/a/c/"));
    }
    if (failures) {
        printf("%d FAILURES\n", failures);
        return 1;
    }
    printf("oracle selftest: all checks passed\n");
    return 0;
}
