// arena_layout_test.cpp — CPU test of the symmetric arena's placement logic (ChunkLayout).
// Properties: live blocks never overlap and stay inside the chunk; the placement is a pure
// function of the call sequence (two instances fed the same sequence agree — what keeps the
// ranks symmetric); releasing everything, in any order, returns the chunk to its empty state;
// recycled blocks are reported as such (the runtime clears them), untouched ones are not.
#include <stdio.h>
#include <stdlib.h>

#include <map>
#include <vector>

#include "../../dlnetbench_b200/csrc/arena_layout.hpp"

#define CHECK(c)                                                                  \
    do {                                                                          \
        if (!(c)) { printf("FAILED %s:%d: %s\n", __FILE__, __LINE__, #c); return 1; } \
    } while (0)

int main() {
    using dlnb::ChunkLayout;
    {   // bump, reuse, coalescing
        ChunkLayout L(10240);
        size_t a, b, c, d;
        bool r;
        CHECK(L.take(1024, &a, &r) && a == 0 && !r);
        CHECK(L.take(2048, &b, &r) && b == 1024 && !r);
        CHECK(L.take(1024, &c, &r) && c == 3072 && !r);
        L.give_back(b, 2048);
        CHECK(L.bytes_free_listed() == 2048 && L.high_water() == 4096);
        CHECK(L.take(1024, &d, &r) && d == 1024 && r);           // first fit inside the hole, reported as recycled
        CHECK(L.take(2048, &b, &r) && b == 4096 && !r);           // does not fit the 1 KiB rest: untouched tail
        L.give_back(d, 1024);
        CHECK(L.bytes_free_listed() == 2048);                     // merged with the rest of the hole
        L.give_back(b, 2048);                                     // trailing: lowers the mark
        CHECK(L.high_water() == 4096 && L.bytes_free_listed() == 2048);
        L.give_back(c, 1024);                                     // now everything above `a` is free
        CHECK(L.high_water() == 1024 && L.bytes_free_listed() == 0);
        CHECK(L.take(1024, &c, &r) && c == 1024 && r);            // the tail below the old mark has held data
        L.give_back(c, 1024);
        L.give_back(a, 1024);
        CHECK(L.high_water() == 0 && L.bytes_free_listed() == 0);
        CHECK(!L.take(10241, &a, &r));
        CHECK(L.take(10240, &a, &r) && a == 0);
    }
    {   // randomized: two instances in lock step, a model of the live blocks
        ChunkLayout A(1u << 20), B(1u << 20);
        std::map<size_t, size_t> live; // off -> size
        std::vector<size_t> offs;
        unsigned s = 12345;
        auto rnd = [&] { s = s * 1664525u + 1013904223u; return s >> 8; };
        for (int step = 0; step < 20000; ++step) {
            if (live.empty() || rnd() % 3) {
                size_t need = (1 + rnd() % 24) * 1024, oa = 0, ob = 0;
                bool ra = false, rb = false;
                bool ga = A.take(need, &oa, &ra), gb = B.take(need, &ob, &rb);
                CHECK(ga == gb);
                if (!ga) continue;
                CHECK(oa == ob && ra == rb && oa % 1024 == 0 && oa + need <= A.size());
                auto nx = live.lower_bound(oa);
                CHECK(nx == live.end() || oa + need <= nx->first);
                if (nx != live.begin()) { auto pv = std::prev(nx); CHECK(pv->first + pv->second <= oa); }
                live[oa] = need;
                offs.push_back(oa);
            } else {
                size_t k = rnd() % offs.size();
                size_t off = offs[k];
                offs[k] = offs.back();
                offs.pop_back();
                A.give_back(off, live[off]);
                B.give_back(off, live[off]);
                live.erase(off);
            }
            size_t used = 0;
            for (auto &kv : live) used += kv.second;
            CHECK(used + A.bytes_free_listed() == A.high_water());
        }
        for (size_t off : offs) { A.give_back(off, live[off]); }
        CHECK(A.high_water() == 0 && A.bytes_free_listed() == 0);
    }
    printf("OK\n");
    return 0;
}
