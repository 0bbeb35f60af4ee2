#include <cstdint>
__global__ void k(const float4* in, float4* out, int n) {
    unsigned long long a0 = 0, a1 = 0;
    for (int i = threadIdx.x; i < n; i += 32) {
        const ulonglong2 v = *reinterpret_cast<const ulonglong2*>(in + i);
        asm("add.rn.f32x2 %0, %0, %1;" : "+l"(a0) : "l"(v.x));
        asm("add.rn.f32x2 %0, %0, %1;" : "+l"(a1) : "l"(v.y));
    }
    ulonglong2 r; r.x = a0; r.y = a1;
    *reinterpret_cast<ulonglong2*>(out + threadIdx.x) = r;
}
__global__ void k2(const float* in, float* out) {
    float a = in[0], b = in[1], c = in[2];
    float m;
    asm("min.f32 %0, %1, %2, %3;" : "=f"(m) : "f"(a), "f"(b), "f"(c));
    out[0] = m;
}
