/* TEST INFRASTRUCTURE ONLY -- CPU restatement ("twin") of the reference's traversal hot path.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load this library.
 * It is never linked into, imported by, or executed from the product (librtb.so has no CPU fallback).
 *
 * Parity status: PINNED against the reference itself run in the build container (oracle/_ref/libref_oracle.so =
 * the unmodified reference + oracle/ref_harness.cpp): tests/test_oracle.py compares every function below with the
 * reference's own BVH::intersect / ManyLightsDistribution::pdf / <random> outputs, and against the committed
 * golden vectors under tests/golden/ (generated from the reference by tests/golden/make_golden.py).
 * The reference ships no tests or golden vectors of its own.
 *
 * All file:line citations are relative to /root/reference.
 */
#ifndef RTB_TWIN_H_
#define RTB_TWIN_H_
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* flat BVH in REFERENCE node order (struct Node, src/core/bvh.h:10-17), as include/rtb.h describes it */
typedef struct twin_bvh {
    uint32_t n_nodes;
    const float *aabb;      /* n_nodes x 6 : min xyz, max xyz */
    const int32_t *links;   /* n_nodes x 5 : left, right, split_dim, first_primitive_id, primitive_count (-1 = invalidKey) */
    uint32_t n_tris;
    const float *tri_pos;   /* n_tris x 9 : p0, U, V (Primitive::position, src/geometry/primitive.h:101) */
} twin_bvh;

/* BVH::intersect (src/core/bvh.cpp:239-243) for a batch; rays6 = origin + un-normalised direction.
 * id = -1 on a miss.  counters2 (may be NULL) receives {ray/box tests, ray/triangle tests} summed over the batch.
 * Returns seconds spent in the loop.  threads <= 0: all OpenMP threads. */
double twin_closest_hit(const twin_bvh *bvh, const float *rays6, size_t n, int32_t *id, float *t, float *u, float *v,
                        uint8_t *inside, uint64_t *counters2, int threads);

/* ManyLightsDistribution::pdf (src/utils/random.cpp:179-188) over the light BVH, for (point, direction) pairs */
void twin_light_pdf(const twin_bvh *light_bvh, const float *rays6, size_t n, float *pdf);

/* libstdc++ <random> restatement (src/utils/random.h:12-17; bits/random.h, bits/random.tcc) */
uint32_t twin_minstd_first(uint32_t seed);
float twin_uniform_first(uint32_t seed, float a, float b);
void twin_normal_seq(uint32_t seed, int n, float *out);

/* resolve pass of Scene::render (src/core/scene.cpp:54-63): sums -> RGB8 */
void twin_resolve(const float *accum, size_t n_values, int samples, uint8_t *rgb8);

#ifdef __cplusplus
}
#endif
#endif
