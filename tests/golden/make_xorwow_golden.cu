// Generates tests/golden/xorwow_golden.txt from cuRAND's OWN implementation
// (toolkit header curand_kernel.h with its host-side skip-ahead tables), run on
// the CPU:  nvcc -o /tmp/mkx make_xorwow_golden.cu && /tmp/mkx > xorwow_golden.txt
// Each line: seed subsequence offset d v0 v1 v2 v3 v4 then 4 uniforms (hex bits).
#define QUALIFIERS static inline __host__ __device__
#include <curand_kernel.h>
#include <cstdio>
#include <cstring>
int main() {
    const unsigned long long seeds[] = {0ull, 1234ull, 0xdeadbeefcafef00dull};
    const unsigned long long subs[] = {0, 1, 2, 7, 479, 599, 1199, 2132, 4031};
    const unsigned long long offs[] = {0, 1, 31, 32, 639, 1599, 3199, 6047};
    for (auto s : seeds) for (auto q : subs) for (auto o : offs) {
        curandStateXORWOW_t st;
        curand_init(s, q, o, &st);
        printf("%llu %llu %llu %u %u %u %u %u %u", s, q, o, st.d, st.v[0], st.v[1], st.v[2], st.v[3], st.v[4]);
        for (int k = 0; k < 4; ++k) { float u = curand_uniform(&st); unsigned b; memcpy(&b, &u, 4); printf(" %08x", b); }
        printf("\n");
    }
    return 0;
}
