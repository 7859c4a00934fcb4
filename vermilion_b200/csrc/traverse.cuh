// traverse.cuh — device-side closest-hit traversal and the RayCast restatement.
//
// Replaces, per ray: BVH::getIntersection (reference core/accelerators/bvh.cpp:47-145),
// BBox::intersect (bbox.cpp:70-83), Triangle::getIntersection (triangle.cpp:4-54),
// Triangle::getNormal (triangle.cpp:56-87), sphereIntersect and MeshEngine::RayCast
// (core/engines/meshEngine.cpp:182-194, 239-509).
//
// Parity rules: the triangle test and everything that produces an OUTPUT value (t, barycentrics,
// normals, sphere solves) uses the reference's exact fp32/fp64 operation order with no FMA
// contraction (file is compiled -fmad=false). Box tests only have to be conservative, so they use
// FMAs on the quantised boxes. Ties (equal t) keep the first triangle visited, as bvh.cpp:90 does;
// the visiting order differs from the reference's tree, so exact ties may resolve differently.
#pragma once
#include "common.cuh"

namespace vmn {

#define VMN_T_MISS 999999999.f // bvh.cpp:48
#ifndef VMN_SH_STACK
#define VMN_SH_STACK 12        // traversal stack entries per thread kept in shared memory
#endif
#ifndef VMN_LOCAL_STACK
#define VMN_LOCAL_STACK (VMN_STACK_CAP - VMN_SH_STACK) // overflow entries in local memory: 128 in all (kTravStackCap, checked against the tree at commit)
#endif
#ifndef VMN_HIT8
#define VMN_HIT8 2 // 1: one hit bit per child, expanded once per node (+2.7 % on k_extend, C2); 2: the bits come from the sign of tf - tn
#endif
#ifndef VMN_SIGNED_LOADS
#define VMN_SIGNED_LOADS 0 // 1: near / far plane words fetched from sign-dependent addresses (six LDG.64) instead of three LDG.128 + 12 SEL: k_extend +2.7 % (profiles/r01aj_ab_nodetest.txt)
#endif
#ifndef VMN_PREFETCH
#define VMN_PREFETCH 0
#endif
#ifndef VMN_PREFETCH_RAYS
#define VMN_PREFETCH_RAYS 0   // prefetch a claimed chunk of the ray queue into L2 (experiment)
#endif
#ifndef VMN_PREFETCH_TRIS
#define VMN_PREFETCH_TRIS 0   // prefetch the first pending leaf triangle right after the node test (experiment)
#endif
#ifndef VMN_COOP_ATOMIC
#define VMN_COOP_ATOMIC 1     // fold pooled triangle results with a shared 64-bit atomicMin per hit instead of a per-owner loop
#endif
#define VMN_WSCRATCH 96       // words of shared scratch per warp (trav_tris_coop)
#ifndef VMN_TRI_DEFER
#define VMN_TRI_DEFER 16      // > 0: run the pooled triangle phase only once this many (ray, triangle) pairs are pending in the warp (or no
                              // lane can advance otherwise); lanes with untested triangles wait, so every ray's own sequence is unchanged
#endif
#ifndef VMN_COOP
#define VMN_COOP 1            // warp-pooled triangle tests (trav_tris_coop); 0 = per-lane loop
#endif

static_assert(VMN_SH_STACK + VMN_LOCAL_STACK == (int)kTravStackCap, "shared + local stack entries must add up to kTravStackCap");

#ifndef VMN_DEFER_STORE
#define VMN_DEFER_STORE 1 // store a finished ray's result with the next refill (several idle lanes) instead of at once: C3 1302 -> 1407 Mrays/s,
                          // C2 unchanged (profiles/r02d_ab_bisect.txt). Tried with it and dropped: prefetch.global.L2 of every hit leaf's first
                          // triangle (C3 +-0) and of the hit sibling nodes that are visited later (C3 -10 %; both +2.3 % on C2's k_extend)
#endif
#ifndef VMN_ABANDON
#define VMN_ABANDON 0     // stack overflow: 1 = abandon the ray after reporting, 0 = report and go on without the dropped siblings. The early
                          // return of 1 costs the node loop its load scheduling: k_extend +39 % on C2, C3 1390 -> 1020 Mrays/s
                          // (profiles/r02d_ab_bisect.txt). The branch is unreachable in the product build (commit refuses trees deeper than
                          // the 128-entry stack); the 3-entry `tinystack` test build, where it is reached, is compiled with 1.
#endif

struct TravCounters { unsigned long long nodes, tris; };

// Queue traffic (rays in, hits out, rays appended by the shading stage) is touched once per launch and is far larger than the L2: it is
// issued with the evict-first cache operators (ld.global.cs / st.global.cs), so that it does not push the nodes and triangles every ray
// re-reads out of the L2 (on C2, 1.07 GB of the 2.40 GB of DRAM traffic per k_extend launch was scene re-fetch, profiles/r01ak).
#ifndef VMN_STREAM_HINTS
#define VMN_STREAM_HINTS 1
#endif
template <class T> __device__ __forceinline__ T ld_stream(const T* p) {
#if VMN_STREAM_HINTS
    return __ldcs(p);
#else
    return __ldg(p);
#endif
}
template <class T> __device__ __forceinline__ void st_stream(T* p, T v) {
#if VMN_STREAM_HINTS
    __stcs(p, v);
#else
    *p = v;
#endif
}

// Triangle::getIntersection (triangle.cpp:4-54): Moeller-Trumbore, two-sided.
__device__ __forceinline__ bool tri_intersect(const TriRec& T, V3 o, V3 d, float* t) {
    V3 v0(T.r0.x, T.r0.y, T.r0.z), e1(T.r1.x, T.r1.y, T.r1.z), e2(T.r2.x, T.r2.y, T.r2.z);
    V3 pvec = cross(d, e2);
    float det = dot(e1, pvec);
    // :25 compares (double)det with +-1e-8; (float)1e-8 = 0x1.5798eep-27 < 1e-8, so for a float det that is |det| <= (float)1e-8
    if (fabsf(det) <= 0x1.5798eep-27f) return false;
    float inv_det = 1 / det;
    V3 tvec = o - v0;
    float u = dot(tvec, pvec) * inv_det;
    if (u < 0 || u > 1) return false;
    V3 qvec = cross(tvec, e1);
    float v = dot(d, qvec) * inv_det;
    if (v < 0 || u + v > 1) return false;
    float dist = dot(e2, qvec) * inv_det;
    if (dist > 0.0f) { *t = dist; return true; }
    return false;
}

__device__ __forceinline__ TriRec load_tri(const TriRec* __restrict__ p) {
    const float4* q = reinterpret_cast<const float4*>(p);
    TriRec t; t.r0 = __ldg(q); t.r1 = __ldg(q + 1); t.r2 = __ldg(q + 2);
    return t;
}
__device__ __forceinline__ TriRec load_tri(const LeafTri* __restrict__ p) {
    const float4* q = reinterpret_cast<const float4*>(p);
    TriRec t; t.r0 = __ldg(q); t.r1 = __ldg(q + 1); t.r2 = __ldg(q + 2);
    return t;
}
// original triangle id of leaf-order record k (or VMN_MISS)
__device__ __forceinline__ uint32_t leaf_original_id(const LeafTri* __restrict__ tris, uint32_t k) {
    return k == VMN_MISS ? VMN_MISS : __float_as_uint(__ldg(&reinterpret_cast<const float4*>(tris + k)->w));
}

__device__ __forceinline__ bool ray_is_nan(V3 o, V3 d) {
    return !(o.x == o.x && o.y == o.y && o.z == o.z && d.x == d.x && d.y == d.y && d.z == d.z);
}

// 1 / d for the slab tests only (never for an output value): MUFU.RCP, 1 ulp. Both planes of an axis use the same reciprocal, so its error
// moves an axis interval as a whole by a relative 2^-23 — the same size as the rounding of (p - o) * idx that the slab margins already absorb.
__device__ __forceinline__ float safe_inv(float x) {
    const float eps = 1e-20f;
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(fabsf(x) > eps ? x : copysignf(eps, x)));
    return r;
}

// ---- compressed BVH8 closest hit --------------------------------------------------------------
// Traversal state of one ray, kept in registers so that a lane can finish one ray and start the next inside the
// same persistent warp (see k_extend). sstack: this thread's slice of the shared-memory stack, entries `sstride`
// uint2 apart; deeper entries spill to a small local-memory array.
struct Trav {
    V3 o, d;
    float idx, idy, idz;
    uint32_t oct_inv;
    uint2 ng, tg;   // node group: (first child index, hit bits 31..24 | imask 7..0); triangle group: (base, hit bits 23..0)
    int sp;
    float best_t;
    uint32_t best;  // index into sc.tris
};

// Quantised byte J of word w as the float 1 + q * 2^-15: one PRMT builds 0x3F80qq00 (q in mantissa bits 8..15). The plane
// distance q*a + b is then evaluated as fma(1 + q*2^-15, a*2^15, b - a*2^15): one PRMT + one FFMA per plane. (I2F.U8 runs on the
// quarter-rate XU pipe and was 33 % of all stall samples in the first profile; the 2^23-bias form needed an extra FADD per plane.)
// Rounding of b - a*2^15 costs at most 2^-9 of a quantisation cell; trav_node() widens every slab by 2^-8 cell to stay conservative.
// `one` must be 0x3F800000 held in a REGISTER (see opaque_one): PRMT takes a single immediate, and with the constant known ptxas spends
// it on 0x3F800000 and re-materialises the selector from a uniform register before every PRMT (one extra IMAD.U32 per plane, 48 per node).
template <int J>
__device__ __forceinline__ float byte_to_unit(uint32_t w, uint32_t one) {
    return __uint_as_float(__byte_perm(w, one, 0x7604u | (J << 4)));
}
// 0x3F800000 that the compiler cannot constant-fold (n_tris < 2^31 always)
__device__ __forceinline__ uint32_t opaque_one(const SceneDev& sc) { return 0x3F800000u | (sc.n_tris >> 31); }

__device__ __forceinline__ bool trav_init(Trav& s, const SceneDev& sc, V3 o, V3 d) {
    s.o = o; s.d = d;
    s.best_t = VMN_T_MISS; s.best = VMN_MISS;
    s.sp = 0;
    s.ng = make_uint2(0u, 0x80000000u); // root: one pending hit at bit 31, imask 0
    s.tg = make_uint2(0u, 0u);
    if (sc.n_tris == 0 || ray_is_nan(o, d)) return false; // nothing to traverse: miss
    s.idx = safe_inv(d.x); s.idy = safe_inv(d.y); s.idz = safe_inv(d.z);
    s.oct_inv = (d.x >= 0.f ? 4u : 0u) | (d.y >= 0.f ? 2u : 0u) | (d.z >= 0.f ? 1u : 0u);
    return true;
}

template <int J>
__device__ __forceinline__ uint32_t child_hit(uint32_t metaw, uint32_t nxw, uint32_t nyw, uint32_t nzw, uint32_t fxw, uint32_t fyw, uint32_t fzw,
                                              float ax, float ay, float az, float nbx, float nby, float nbz, float fbx, float fby, float fbz,
                                              float best_t, uint32_t oct_inv, uint32_t one) {
    const float tnx = __fmaf_rn(byte_to_unit<J>(nxw, one), ax, nbx);
    const float tny = __fmaf_rn(byte_to_unit<J>(nyw, one), ay, nby);
    const float tnz = __fmaf_rn(byte_to_unit<J>(nzw, one), az, nbz);
    const float tfx = __fmaf_rn(byte_to_unit<J>(fxw, one), ax, fbx);
    const float tfy = __fmaf_rn(byte_to_unit<J>(fyw, one), ay, fby);
    const float tfz = __fmaf_rn(byte_to_unit<J>(fzw, one), az, fbz);
    const float tn = fmaxf(fmaxf(tnx, tny), fmaxf(tnz, 0.f));
    const float tf = fminf(fminf(tfx, tfy), fminf(tfz, best_t));
#if VMN_HIT8 == 2
    return __float_as_uint(tf - tn); // sign bit set = miss (no NaNs here: safe_inv keeps every slab distance finite); see trav_node
#elif VMN_HIT8
    return tn <= tf ? 1u : 0u; // trav_node expands the 8 hit bits once per node
#endif
    if (tn <= tf) {
        const uint32_t m = (metaw >> (8 * J)) & 0xFFu;
        const uint32_t inner = ((m & 0x18u) == 0x18u) ? 7u : 0u;
        const uint32_t bitidx = (m ^ (oct_inv & inner)) & 0x1Fu;
        return (m >> 5) << bitidx;
    }
    return 0u;
}

// A traversal step has three phases, so that the triangle tests of a whole warp can be pooled (trav_tris_coop):
//   trav_node : visit the next pending node (8 child boxes) or turn a popped triangle group into pending triangles
//   trav_tris : test the pending triangles
//   trav_pop  : pop the next node group if none is pending; false when the ray is finished
template <bool COUNT>
__device__ __forceinline__ void trav_node(Trav& s, const SceneDev& sc, uint2* sstack, uint32_t sstride, uint2* lstack, TravCounters* cnt) {
    const uint4* __restrict__ nodes = reinterpret_cast<const uint4*>(sc.nodes);
    if (s.ng.y & 0xFF000000u) {
        const uint32_t bit = 31u - __clz(s.ng.y);
        const uint32_t slot = (bit - 24u) ^ s.oct_inv;
        s.ng.y &= ~(1u << bit);
        const uint32_t rel = __popc(s.ng.y & 0xFFu & ((1u << slot) - 1u));
        const uint32_t nidx = s.ng.x + rel;
        if (s.ng.y & 0xFF000000u) { // siblings still pending
            // one entry per wide-tree level at most; vmn_scene_commit refuses a tree deeper than the stack (kTravStackCap), so the last
            // branch is unreachable for a committed scene — if it ever runs, the dropped push is REPORTED (VMN_ERR_STATE), never silent
            if (s.sp < VMN_SH_STACK) { sstack[s.sp * sstride] = s.ng; ++s.sp; }
            else if (s.sp < VMN_SH_STACK + VMN_LOCAL_STACK) { lstack[s.sp - VMN_SH_STACK] = s.ng; ++s.sp; }
            else { // report, and abandon this ray: nothing pending, nothing stacked -> trav_pop ends it
                atomicOr(sc.err, VMN_DEVERR_STACK);
#if VMN_ABANDON
                s.ng = make_uint2(0u, 0u); s.tg = make_uint2(0u, 0u); s.sp = 0;
                return;
#endif
            }
        }
        VMN_CHECK(sc.err, nidx < sc.n_wide && s.sp >= 0 && s.sp <= (int)kTravStackCap);
        const uint4* np = nodes + (uint64_t)kNodeStride16 * nidx;
#if VMN_SIGNED_LOADS
        // The near plane of an axis is the low plane for a positive direction and the high plane for a negative one. Instead of loading
        // all six plane groups and swapping them with 12 SELs (ALU pipe, the one k_extend saturates), each 8-byte group is fetched from
        // the address the direction's SIGN BIT picks: x low / high at +32 / +56, y at +40 / +64, z at +48 / +72, so near = low + 24 * sign
        // and far = high - 24 * sign. The sign comes from IMAD.HI and the address from IMAD.WIDE (FMA pipe). Using the sign bit rather
        // than d < 0 also pairs -0.0 with the negative reciprocal safe_inv gives it.
        const uint4 q0 = __ldg(np), q1 = __ldg(np + 1);
        const char* nb = reinterpret_cast<const char*>(np);
        const uint32_t sgx = __umulhi(__float_as_uint(s.d.x), 2u), sgy = __umulhi(__float_as_uint(s.d.y), 2u), sgz = __umulhi(__float_as_uint(s.d.z), 2u);
        const uint2 pnx = __ldg(reinterpret_cast<const uint2*>(nb + 32 + (uint64_t)sgx * 24u));
        const uint2 pny = __ldg(reinterpret_cast<const uint2*>(nb + 40 + (uint64_t)sgy * 24u));
        const uint2 pnz = __ldg(reinterpret_cast<const uint2*>(nb + 48 + (uint64_t)sgz * 24u));
        const uint2 pfx = __ldg(reinterpret_cast<const uint2*>(nb + 56 + (int64_t)(int32_t)sgx * -24));
        const uint2 pfy = __ldg(reinterpret_cast<const uint2*>(nb + 64 + (int64_t)(int32_t)sgy * -24));
        const uint2 pfz = __ldg(reinterpret_cast<const uint2*>(nb + 72 + (int64_t)(int32_t)sgz * -24));
#else
        const uint4 q0 = __ldg(np), q1 = __ldg(np + 1), q2 = __ldg(np + 2), q3 = __ldg(np + 3), q4 = __ldg(np + 4);
#endif
        if (COUNT) ++cnt->nodes;
        // per-axis: t(q) = q * a + b with a = 2^(e-127) / d, b = (p - o) / d, evaluated as fma(1 + q*2^-15, A, B), A = a * 2^15,
        // B = b - A, widened by |a| * 2^-8 (see byte_to_unit)
        const float ax = __uint_as_float((q0.w & 0xFFu) << 23) * s.idx;
        const float ay = __uint_as_float(((q0.w >> 8) & 0xFFu) << 23) * s.idy;
        const float az = __uint_as_float(((q0.w >> 16) & 0xFFu) << 23) * s.idz;
        const float Ax = ax * 32768.f, Ay = ay * 32768.f, Az = az * 32768.f;
        const float bx = (__uint_as_float(q0.x) - s.o.x) * s.idx - Ax;
        const float by = (__uint_as_float(q0.y) - s.o.y) * s.idy - Ay;
        const float bz = (__uint_as_float(q0.z) - s.o.z) * s.idz - Az;
        const float mx = fabsf(ax) * 0.00390625f, my = fabsf(ay) * 0.00390625f, mz = fabsf(az) * 0.00390625f;
        const float nbx = bx - mx, nby = by - my, nbz = bz - mz, fbx = bx + mx, fby = by + my, fbz = bz + mz;
        // near/far byte planes per axis (two words = 8 children each)
#if VMN_SIGNED_LOADS
        const uint32_t nx0 = pnx.x, nx1 = pnx.y, fx0 = pfx.x, fx1 = pfx.y;
        const uint32_t ny0 = pny.x, ny1 = pny.y, fy0 = pfy.x, fy1 = pfy.y;
        const uint32_t nz0 = pnz.x, nz1 = pnz.y, fz0 = pfz.x, fz1 = pfz.y;
#else
        // the SIGN BIT, not d < 0: safe_inv gives -0.0 a negative reciprocal, so its near plane is the high one (with d < 0 such a ray
        // missed every box: tests/test_gpu_parity.py::test_signed_zero_direction_components_match_oracle)
        const bool ngx = __float_as_int(s.d.x) < 0, ngy = __float_as_int(s.d.y) < 0, ngz = __float_as_int(s.d.z) < 0;
        const uint32_t nx0 = ngx ? q3.z : q2.x, nx1 = ngx ? q3.w : q2.y, fx0 = ngx ? q2.x : q3.z, fx1 = ngx ? q2.y : q3.w;
        const uint32_t ny0 = ngy ? q4.x : q2.z, ny1 = ngy ? q4.y : q2.w, fy0 = ngy ? q2.z : q4.x, fy1 = ngy ? q2.w : q4.y;
        const uint32_t nz0 = ngz ? q4.z : q3.x, nz1 = ngz ? q4.w : q3.y, fz0 = ngz ? q3.x : q4.z, fz1 = ngz ? q3.y : q4.w;
#endif
        uint32_t hitmask = 0;
        const uint32_t one = opaque_one(sc);
#if VMN_HIT8
        // one hit bit per child slot, then a single expansion per node: internal hits are moved to the octant-ordered positions
        // 24 + (slot ^ oct_inv) by three conditional bit swaps; leaf hits OR in their unary triangle masks (usually 0-2 leaves)
#if VMN_HIT8 == 2
        // the hit test tn <= tf is the sign of tf - tn (FADD, FMA pipe); the eight signs are shifted into one register by funnel shifts,
        // child 7 first so that child j lands on bit j: 8 FADD + 8 SHF instead of 8 FSETP + 8 SEL + 4 IADD3 on the ALU pipe
        uint32_t miss8 = 0;
#define VMN_CH(J, MW, S) miss8 = __funnelshift_l(child_hit<J>(MW, nx##S, ny##S, nz##S, fx##S, fy##S, fz##S, Ax, Ay, Az, nbx, nby, nbz, fbx, fby, fbz, s.best_t, s.oct_inv, one), miss8, 1)
        VMN_CH(3, q1.w, 1); VMN_CH(2, q1.w, 1); VMN_CH(1, q1.w, 1); VMN_CH(0, q1.w, 1);
        VMN_CH(3, q1.z, 0); VMN_CH(2, q1.z, 0); VMN_CH(1, q1.z, 0); VMN_CH(0, q1.z, 0);
#undef VMN_CH
        const uint32_t hit8 = ~miss8 & 0xFFu;
#else
        uint32_t hit8 = 0;
#define VMN_CH(J, MW, S) hit8 |= child_hit<J>(MW, nx##S, ny##S, nz##S, fx##S, fy##S, fz##S, Ax, Ay, Az, nbx, nby, nbz, fbx, fby, fbz, s.best_t, s.oct_inv, one) << (4 * S + J)
        VMN_CH(0, q1.z, 0); VMN_CH(1, q1.z, 0); VMN_CH(2, q1.z, 0); VMN_CH(3, q1.z, 0);
        VMN_CH(0, q1.w, 1); VMN_CH(1, q1.w, 1); VMN_CH(2, q1.w, 1); VMN_CH(3, q1.w, 1);
#undef VMN_CH
#endif
        {
            const uint32_t imask = q0.w >> 24;
            uint32_t x = hit8 & imask;
            if (s.oct_inv & 1u) x = ((x & 0x55u) << 1) | ((x >> 1) & 0x55u);
            if (s.oct_inv & 2u) x = ((x & 0x33u) << 2) | ((x >> 2) & 0x33u);
            if (s.oct_inv & 4u) x = ((x & 0x0Fu) << 4) | ((x >> 4) & 0x0Fu);
            hitmask = x << 24;
            uint32_t hl = hit8 & ~imask;
            while (hl) {
                const uint32_t j = __ffs(hl) - 1u;
                hl &= hl - 1u;
                const uint32_t m = __byte_perm(q1.z, q1.w, j) & 0xFFu; // byte j of the 8 meta bytes
                hitmask |= (m >> 5) << (m & 31u);
            }
        }
#else
#define VMN_CH(J, MW, S) hitmask |= child_hit<J>(MW, nx##S, ny##S, nz##S, fx##S, fy##S, fz##S, Ax, Ay, Az, nbx, nby, nbz, fbx, fby, fbz, s.best_t, s.oct_inv, one)
        VMN_CH(0, q1.z, 0); VMN_CH(1, q1.z, 0); VMN_CH(2, q1.z, 0); VMN_CH(3, q1.z, 0);
        VMN_CH(0, q1.w, 1); VMN_CH(1, q1.w, 1); VMN_CH(2, q1.w, 1); VMN_CH(3, q1.w, 1);
#undef VMN_CH
#endif
        s.ng = make_uint2(q1.x, (hitmask & 0xFF000000u) | (q0.w >> 24));
        s.tg = make_uint2(q1.y, hitmask & 0x00FFFFFFu);
#if VMN_PREFETCH_TRIS
        if (s.tg.y) asm volatile("prefetch.global.L1 [%0];" ::"l"(sc.tris + s.tg.x + (__ffs(s.tg.y) - 1u)));
#endif
#if VMN_PREFETCH
        // the nearest internal child hit is the next node this ray visits (unless a triangle test culls it): start pulling it into L2
        // now, so that its DRAM latency overlaps the triangle phase instead of following it
        if (s.ng.y & 0xFF000000u) {
            const uint32_t nslot = ((31u - __clz(s.ng.y)) - 24u) ^ s.oct_inv;
            const uint4* nx = nodes + (uint64_t)kNodeStride16 * (s.ng.x + __popc(s.ng.y & 0xFFu & ((1u << nslot) - 1u)));
            asm volatile("prefetch.global.L2 [%0];" ::"l"(nx));
#if VMN_PREFETCH > 1
            asm volatile("prefetch.global.L2 [%0];" ::"l"(nx + 4));
#endif
        }
#endif
    }
    // (no else: the stack only ever holds groups with pending hit bits, and trav_pop refills ng before the next call)
}

// per-lane triangle loop (single-ray form)
template <bool COUNT>
__device__ __forceinline__ void trav_tris(Trav& s, const SceneDev& sc, TravCounters* cnt) {
    while (s.tg.y) {
        const uint32_t i = __ffs(s.tg.y) - 1u;
        s.tg.y &= s.tg.y - 1u;
        const uint32_t ti = s.tg.x + i;
        VMN_CHECK(sc.err, ti < sc.n_tris);
        const TriRec T = load_tri(sc.tris + ti);
        if (COUNT) ++cnt->tris;
        float ct;
        if (tri_intersect(T, s.o, s.d, &ct) && ct < s.best_t) { s.best_t = ct; s.best = ti; }
    }
}

__device__ __forceinline__ bool trav_pop(Trav& s, const uint2* sstack, uint32_t sstride, const uint2* lstack) {
    if ((s.ng.y & 0xFF000000u) == 0u) {
        if (s.sp == 0) return false;
        --s.sp;
        s.ng = (s.sp < VMN_SH_STACK) ? sstack[s.sp * sstride] : lstack[s.sp - VMN_SH_STACK];
    }
    return true;
}

// Warp-pooled triangle tests. After a node visit only ~3 of 32 lanes have triangles pending, and a per-lane loop makes
// the whole warp wait for the lane with the most (the first ncu source view: 43 % of k_extend's issue slots at 3.3 active
// threads). Here the warp's pending (ray, triangle) pairs are numbered by a prefix sum and dealt one per lane: lane j takes
// pair j, fetches the owner's ray by shuffle, tests, and posts t to shared memory; each owner then folds its results in
// triangle order with the reference's strict `<` (bvh.cpp:90), so the outcome is identical to the per-lane loop.
// wscratch: VMN_WSCRATCH words of shared memory per warp. Must be called by all 32 lanes.
template <bool COUNT>
__device__ __forceinline__ void trav_tris_coop(Trav& s, bool active, const SceneDev& sc, uint32_t* wscratch, TravCounters* cnt) {
    const uint32_t FULL = 0xffffffffu;
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t bits = active ? s.tg.y : 0u;
    const uint32_t n = __popc(bits);
    const uint32_t total = __reduce_add_sync(FULL, n);
    if (total == 0u) return;
#if VMN_TRI_DEFER
    // postpone a thin pool while some active lane can still advance without it (warp-uniform decision)
    if (total < (uint32_t)VMN_TRI_DEFER && __ballot_sync(FULL, active && n == 0u) != 0u) return;
#endif
    uint32_t incl = n;
#pragma unroll
    for (int dlt = 1; dlt < 32; dlt <<= 1) {
        const uint32_t up = __shfl_up_sync(FULL, incl, dlt);
        if ((int)lane >= dlt) incl += up;
    }
    const uint32_t excl = incl - n;
    uint32_t* w_owner = wscratch;
#if VMN_COOP_ATOMIC
    // results are folded with one 64-bit shared atomicMin per HIT on the owner's key (t bits << 32 | leaf index): positive floats order
    // as their bit patterns, the smaller leaf index wins an exact tie among the pooled triangles (= triangle order), and the owner's
    // current best enters with index 0 so that it wins ties against all of them — the strict `<` of bvh.cpp:90. wscratch: 96 words.
    unsigned long long* w_key = reinterpret_cast<unsigned long long*>(wscratch + 32);
    const unsigned long long key0 = (unsigned long long)__float_as_uint(s.best_t) << 32;
    if (n) w_key[lane] = key0;
#else
    float* w_res = reinterpret_cast<float*>(wscratch + 32);
#endif
    for (uint32_t w0 = 0; w0 < total; w0 += 32u) { // warp-uniform; one trip unless more than 32 pairs are pending
        const bool has = n != 0u && excl < w0 + 32u && incl > w0;
        const uint32_t lo = excl > w0 ? excl : w0;          // first of my pairs inside the window
        const uint32_t p0 = lo - w0;
        if (has) w_owner[p0] = lane;
        const uint32_t heads = __reduce_or_sync(FULL, has ? (1u << p0) : 0u);
        __syncwarp();
        const uint32_t j = w0 + lane;
        const bool work = j < total;
        uint32_t owner = lane;
        if (work) owner = w_owner[31u - __clz(heads & (0xFFFFFFFFu >> (31u - lane)))];
        const uint32_t o_excl = __shfl_sync(FULL, excl, owner), o_bits = __shfl_sync(FULL, bits, owner), o_base = __shfl_sync(FULL, s.tg.x, owner);
        const float ox = __shfl_sync(FULL, s.o.x, owner), oy = __shfl_sync(FULL, s.o.y, owner), oz = __shfl_sync(FULL, s.o.z, owner);
        const float dx = __shfl_sync(FULL, s.d.x, owner), dy = __shfl_sync(FULL, s.d.y, owner), dz = __shfl_sync(FULL, s.d.z, owner);
        if (work) {
            // my triangle = the (j - o_excl)-th set bit of the owner's mask: ranks 0..2 without a loop (an owner rarely pools more than
            // three triangles; the divergent loop over all ranks was 4.8 % of k_extend's instructions at 5 active lanes)
            const uint32_t rank = j - o_excl;
            const uint32_t b1 = o_bits & (o_bits - 1u), b2 = b1 & (b1 - 1u);
            uint32_t b = rank == 0u ? o_bits : (rank == 1u ? b1 : b2);
            for (uint32_t k = rank; k > 2u; --k) b &= b - 1u;
            const uint32_t ti = o_base + (__ffs(b) - 1u);
            VMN_CHECK(sc.err, ti < sc.n_tris && b != 0u && owner < 32u);
            const TriRec T = load_tri(sc.tris + ti);
            if (COUNT) ++cnt->tris;
            float ct;
#if VMN_COOP_ATOMIC
            if (tri_intersect(T, V3(ox, oy, oz), V3(dx, dy, dz), &ct))
                atomicMin(&w_key[owner], ((unsigned long long)__float_as_uint(ct) << 32) | ti);
#else
            w_res[lane] = tri_intersect(T, V3(ox, oy, oz), V3(dx, dy, dz), &ct) ? ct : INFINITY;
#endif
        }
        __syncwarp();
#if !VMN_COOP_ATOMIC
        if (has) {
            uint32_t b = bits;
            for (uint32_t k = lo - excl; k; --k) b &= b - 1u;
            const uint32_t hi = incl < w0 + 32u ? incl : w0 + 32u;
            for (uint32_t q = p0; q < p0 + (hi - lo); ++q) {
                const float ct = w_res[q];
                const uint32_t ti = s.tg.x + (__ffs(b) - 1u);
                b &= b - 1u;
                if (ct < s.best_t) { s.best_t = ct; s.best = ti; }
            }
        }
        __syncwarp();
#endif
    }
#if VMN_COOP_ATOMIC
    if (n) {
        const unsigned long long k = w_key[lane];
        if (k != key0) { s.best_t = __uint_as_float((uint32_t)(k >> 32)); s.best = (uint32_t)k; }
    }
#endif
    s.tg.y = 0u;
}

// single-ray form: returns the index into sc.tris of the closest triangle (or VMN_MISS) and its t
template <bool COUNT>
__device__ __forceinline__ uint32_t trace_wide(const SceneDev& sc, V3 o, V3 d, float* t_out, uint2* sstack, uint32_t sstride,
                                               TravCounters* cnt) {
    Trav s;
    uint2 lstack[VMN_LOCAL_STACK];
    if (trav_init(s, sc, o, d))
        do {
            trav_node<COUNT>(s, sc, sstack, sstride, lstack, cnt);
            trav_tris<COUNT>(s, sc, cnt);
        } while (trav_pop(s, sstack, sstride, lstack));
    *t_out = s.best_t;
    return s.best;
}

// ---- persistent lanes ----------------------------------------------------------------------------------------------
// Every lane of a persistent warp owns one ray at a time. A warp claims CHUNKS of consecutive rays from the global
// counter (one atomicAdd per chunk, by one lane, broadcast with a shuffle) and hands them to idle lanes by ballot/popc
// rank, as soon as at least `refill_min` lanes are idle. The first profile of the one-batch-per-warp version showed
// 3.4 of 32 threads active per instruction: a few long rays kept a warp alive while the other lanes sat finished.
// load(i, &o, &d) reads ray i; store(i, trav) writes its result. wscratch: 64 words of shared memory per warp.
template <bool COUNT, class LoadRay, class StoreHit>
__device__ __forceinline__ void persistent_trace(const SceneDev& sc, uint64_t n, unsigned long long* next, uint32_t chunk, uint32_t refill_min,
                                                 uint2* sstack, uint32_t sstride, uint32_t* wscratch, TravCounters* cnt, LoadRay load,
                                                 StoreHit store, const void* rays_base = nullptr) {
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t FULL = 0xffffffffu;
    Trav s;
    s.tg = make_uint2(0u, 0u);
    uint2 lstack[VMN_LOCAL_STACK];
    bool active = false, exhausted = false;
    // s.sp == -1 marks "this lane's finished ray has not been stored yet": store() runs with the next refill, when several lanes are
    // idle, instead of once per finishing lane (ncu, C3: the dependent loads of store() were 10.7 % of all stall samples at 1.6 active
    // lanes, stalling the 25 lanes that were still traversing). The flag lives in the stack pointer: no extra register.
    s.sp = 0;
    uint64_t ray = 0, wbase = 0;
    uint32_t wleft = 0; // warp-uniform: rays left in the claimed chunk [wbase, wbase + wleft)
    while (true) {
        const uint32_t idle = __ballot_sync(FULL, !active);
        if (idle != 0u && !exhausted && (idle == FULL || (uint32_t)__popc(idle) >= refill_min)) {
            if (wleft == 0u) {
                unsigned long long b = 0;
                if (lane == 0) b = atomicAdd(next, (unsigned long long)chunk);
                b = __shfl_sync(FULL, b, 0);
                if (b >= n) exhausted = true;
                else {
                    wbase = b; wleft = (uint32_t)((n - b < chunk) ? (n - b) : chunk);
#if VMN_PREFETCH_RAYS
                    for (uint32_t k = 4u * lane; k < wleft; k += 128u) { // one 128-byte line = four 32-byte rays
                        const char* a = reinterpret_cast<const char*>(rays_base) + 32ull * (wbase + k);
#if VMN_PREFETCH_RAYS == 2
                        asm volatile("prefetch.global.L1 [%0];" ::"l"(a));
#else
                        asm volatile("prefetch.global.L2 [%0];" ::"l"(a));
#endif
                    }
#endif
                }
            }
            const uint32_t need = (uint32_t)__popc(idle);
            const uint32_t take = need < wleft ? need : wleft;
            if (!active && s.sp < 0) { store(ray, s); s.sp = 0; }
            if (!active) {
                const uint32_t r = (uint32_t)__popc(idle & ((1u << lane) - 1u));
                if (r < take) {
                    V3 o, d;
                    ray = wbase + r;
                    VMN_CHECK(sc.err, ray < n);
                    load(ray, &o, &d);
                    if (trav_init(s, sc, o, d)) active = true; else store(ray, s);
                }
            }
            wbase += take; wleft -= take;
        }
        if (__ballot_sync(FULL, active) == 0u) {
            if (exhausted) break;
            continue;
        }
        if (exhausted && !active && s.sp < 0) { store(ray, s); s.sp = 0; } // no refill will come: store as lanes finish
#if VMN_TRI_DEFER && VMN_COOP
        if (active && s.tg.y == 0u) trav_node<COUNT>(s, sc, sstack, sstride, lstack, cnt);
#else
        if (active) trav_node<COUNT>(s, sc, sstack, sstride, lstack, cnt);
#endif
#if VMN_COOP
        __syncwarp();
        trav_tris_coop<COUNT>(s, active, sc, wscratch, cnt);
#else
        if (active) trav_tris<COUNT>(s, sc, cnt);
#endif
        if (active && s.tg.y == 0u && !trav_pop(s, sstack, sstride, lstack)) {
#if VMN_DEFER_STORE
            s.sp = -1;
#else
            store(ray, s);
#endif
            active = false;
        }
    }
    if (!active && s.sp < 0) store(ray, s);
}

__device__ __forceinline__ uint32_t pick_chunk(uint64_t n, uint32_t warps) {
    uint64_t c = n / ((uint64_t)warps * 4u);
    c = c < 32 ? 32 : (c > 256 ? 256 : c);
    return (uint32_t)(c & ~31ull);
}

// ---- binary LBVH closest hit (baseline / cross-check; stack in local memory) --------------------
template <bool COUNT>
__device__ __forceinline__ uint32_t trace_bvh2(const SceneDev& sc, V3 o, V3 d, float* t_out, TravCounters* cnt) {
    float best_t = VMN_T_MISS;
    uint32_t best = VMN_MISS; // index into sc.tris_sorted
    const uint32_t n = sc.bvh2.n_prims;
    if (n == 0 || ray_is_nan(o, d)) { *t_out = best_t; return best; }
    const int nint = (int)n - 1;
    const float idx = safe_inv(d.x), idy = safe_inv(d.y), idz = safe_inv(d.z);
    int stack[kTravStackCap];
    int sp = 0;
    int node = 0;
    while (true) {
        if (node >= nint) {
            const uint32_t k = (uint32_t)(node - nint);
            VMN_CHECK(sc.err, k < n);
            const TriRec T = load_tri(sc.tris_sorted + k);
            if (COUNT) ++cnt->tris;
            float ct;
            if (tri_intersect(T, o, d, &ct) && ct < best_t) { best_t = ct; best = k; }
            if (sp == 0) break;
            node = stack[--sp];
            continue;
        }
        if (COUNT) ++cnt->nodes;
        VMN_CHECK(sc.err, node >= 0 && node < nint);
        const int2 c = __ldg(&sc.bvh2.child[node]);
        float tn[2]; bool hit[2];
#pragma unroll
        for (int k = 0; k < 2; ++k) {
            const int ci = k ? c.y : c.x;
            const float4 lo = __ldg(&sc.bvh2.bmin[ci]), hi = __ldg(&sc.bvh2.bmax[ci]);
            const float x0 = (lo.x - o.x) * idx, x1 = (hi.x - o.x) * idx;
            const float y0 = (lo.y - o.y) * idy, y1 = (hi.y - o.y) * idy;
            const float z0 = (lo.z - o.z) * idz, z1 = (hi.z - o.z) * idz;
            const float a = fmaxf(fmaxf(fminf(x0, x1), fminf(y0, y1)), fmaxf(fminf(z0, z1), 0.f));
            const float b = fminf(fminf(fmaxf(x0, x1), fmaxf(y0, y1)), fminf(fmaxf(z0, z1), best_t));
            // widen by a few ulp so that rounding in the slab products cannot reject a box the exact test accepts
            hit[k] = a * 0.9999995f <= b * 1.0000005f;
            tn[k] = a;
        }
        if (hit[0] && hit[1]) {
            const bool swap = tn[1] < tn[0];
            const int nearc = swap ? c.y : c.x, farc = swap ? c.x : c.y;
            if (sp < (int)kTravStackCap) stack[sp++] = farc;
            else { atomicOr(sc.err, VMN_DEVERR_STACK); break; } // report, and abandon this ray
            node = nearc;
        } else if (hit[0]) node = c.x;
        else if (hit[1]) node = c.y;
        else {
            if (sp == 0) break;
            node = stack[--sp];
        }
    }
    *t_out = best_t;
    return best;
}

// Barycentric weights (w1, w2) of Triangle::getNormal (triangle.cpp:67-78) for hit point hp.
__device__ __forceinline__ void tri_bary(const TriRec& T, V3 hp, float* w1, float* w2) {
    V3 f0(T.r1.x, T.r1.y, T.r1.z), f1(T.r2.x, T.r2.y, T.r2.z);
    V3 f2 = hp - V3(T.r0.x, T.r0.y, T.r0.z);
    float d00 = dot(f0, f0), d01 = dot(f0, f1), d11 = dot(f1, f1), d20 = dot(f2, f0), d21 = dot(f2, f1);
    float denom = d00 * d11 - d01 * d01;
    *w1 = (d11 * d20 - d01 * d21) / denom;
    *w2 = (d00 * d21 - d01 * d20) / denom;
}

// sphereIntersect (meshEngine.cpp:182-194): fp32 operands, fp64 solve, eps 1e-4. r2 = (double)(rad * rad), precomputed (SphereDev).
__device__ __forceinline__ float sphere_intersect(V3 pos, V3 rot, V3 p, const double r2) {
    V3 op = p - pos;
    double t;
    const double eps = 1e-4;
    double b = (double)dot(op, rot);
    double det = b * b - (double)dot(op, op) + r2;
    if (det < 0) return 0;
    det = sqrt(det);
    return (float)((t = b - det) > eps ? t : ((t = b + det) > eps ? t : 0));
}

struct RayCastOut {
    bool hit, material;
    V3 P, N, Le;
    float dist;
    float uv0, uv1;
    uint32_t tri;
};

// Everything MeshEngine::RayCast does after the BVH query (meshEngine.cpp:364-508).
// bvh_t / tri_id: closest BVH hit (original triangle id or VMN_MISS).
__device__ __forceinline__ RayCastOut raycast_finish(const SceneDev& sc, const SphereDev* __restrict__ sph, V3 o, V3 d, float bvh_t,
                                                     uint32_t tri_id, const bool attrs = true) {
    // attrs = false: the caller ends the path at this hit whatever it is (last segment of a FIXED path) and only reads hit / Le / dist —
    // the triangle's two record gathers, the barycentrics and the normalisation are skipped
    RayCastOut r;
    r.N = V3(0, 0, 0); r.Le = V3(0, 0, 0); r.uv0 = 0.f; r.uv1 = 0.f; r.tri = tri_id;
    r.material = tri_id != VMN_MISS;
    float nearest = r.material ? bvh_t : INFINITY;
    // meshEngine.cpp:377-500: fixed order, accept 0 < ts < nearest. The reference overwrites the normal at every accepted sphere
    // and keeps the emission of the last accepted EMITTER, so only the last accepted sphere's normal survives: it is computed once,
    // after the loop, with the same expression (an IEEE normalize per accept was 11 % of k_shade).
    int last = -1;
    for (uint32_t i = 0; i < sc.n_spheres; ++i) {
        const float4 s0 = __ldg(reinterpret_cast<const float4*>(sph + i));      // centre, flags
        const double2 s1 = __ldg(reinterpret_cast<const double2*>(sph + i) + 1); // r2, (r, pad)
        const float th = sphere_intersect(o, d, V3(s0.x, s0.y, s0.z), s1.x);
        const uint32_t sflags = __float_as_uint(s0.w);
        if (th > 0.f && th < nearest) {
            nearest = th;
            if (sflags & VMN_SPHD_EMITTER) { const float4 e = __ldg(reinterpret_cast<const float4*>(sph + i) + 2); r.Le = V3(e.x, e.y, e.z); }
            last = (int)i;
        }
    }
    if (r.material && attrs) {
        V3 hp = o + d * bvh_t; // bvh.cpp:140
        VMN_CHECK(sc.err, tri_id < sc.n_tris);
        const TriRec T = load_tri(sc.tris_orig + tri_id);
        float w1, w2;
        tri_bary(T, hp, &w1, &w2);
        float w0 = 1 - w1 - w2;
        const float4* sp4 = reinterpret_cast<const float4*>(sc.shade + tri_id);
        const float4 s0 = __ldg(sp4), s1 = __ldg(sp4 + 1), s2 = __ldg(sp4 + 2), s3 = __ldg(sp4 + 3);
        r.uv0 = s0.w * w0 + s2.w * w1 + s3.y * w2; // uv0.x, uv1.x, uv2.x  (triangle.cpp:83)
        r.uv1 = s1.w * w0 + s3.x * w1 + s3.z * w2; // uv0.y, uv1.y, uv2.y
        if (last < 0) {
            V3 n = V3(s0.x, s0.y, s0.z) * w0 + V3(s1.x, s1.y, s1.z) * w1 + V3(s2.x, s2.y, s2.z) * w2; // triangle.cpp:80
            r.N = normalize(-n);                    // triangle.cpp:86, meshEngine.cpp:369
        }
    }
    if (last >= 0 && attrs) {
        const SphereDev s = sph[last];
        V3 nn = normalize(o + (d * nearest) - V3(s.nx, s.ny, s.nz));
        r.N = (s.flags & VMN_SPH_FLIP_NORMAL) ? -nn : nn;
    }
    r.P = o + (d * nearest);
    r.dist = nearest;
    r.hit = nearest < INFINITY;
    return r;
}

} // namespace vmn
