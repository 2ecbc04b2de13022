/* The whole hot path through the C ABI, as a non-Python host would drive it (mirrors
 * Collider.get_collisions, collision/collision.py:130-198 upstream).  Compile-checked by
 * tests/test_host_api.py against include/collision_b200.h; link with -lcollision_b200 -lcudart.
 * All pointers are device pointers the caller allocated (sizes below); `s` is a cudaStream_t. */
#include <stddef.h>
#include <stdint.h>
#include "collision_b200.h"

struct col_buffers {
    const float* coords;   /* n x 4 floats (xyz + pad)            */
    const float* radii;    /* n floats                            */
    uint32_t *codes0, *codes1, *ids0, *ids1; /* n u32 each        */
    float* scene;          /* 2 x 4 floats: {min[4], max[4]}      */
    void* nodes;           /* 16 * (2n-1) bytes: the Node array   */
    float* bounds;         /* 32 * (2n-1) bytes: the Bound array  */
    void* wide;            /* 32 * (n-1) bytes: traversal records */
    uint32_t* flags;       /* 4 * 2n bytes: build scratch          */
    uint32_t* pairs;       /* capacity x 2 u32 rows               */
    uint64_t* total;       /* 1 u64: true number of pairs         */
    void *reduce_scratch, *sort_scratch, *traverse_scratch;
};

/* scratch sizes for n spheres and room for `capacity` pair rows */
void col_example_scratch(uint32_t n, uint64_t capacity, size_t* reduce, size_t* sort, size_t* traverse) {
    *reduce = col_reduce_scratch_bytes();
    *sort = col_sort_scratch_bytes(n, 4, 4);
    *traverse = col_traverse_scratch_bytes(n, capacity);
}

/* enqueue one frame on stream s; returns 0 or the first error (see the header's error codes) */
int col_example_frame(const struct col_buffers* b, uint32_t n, uint64_t capacity, void* s) {
    int rc;
    if ((rc = col_bounds(b->coords, n, COL_F32, 4, b->scene, b->reduce_scratch, s))) return rc;
    if ((rc = col_morton(b->codes0, b->ids0, b->coords, b->scene, COL_F32, n, n, s))) return rc;
    if ((rc = col_sort_pairs(b->codes0, b->codes1, b->ids0, b->ids1, n, 4, 4, /*end_bit=*/30, b->sort_scratch, s)))
        return rc;
    /* tree + leaf boxes + refit + traversal records: one launch */
    if ((rc = col_build_refit_wide(b->codes1, b->ids1, b->nodes, b->wide, b->flags, b->bounds, b->coords, b->radii,
                                   b->scene, COL_F32, n, s)))
        return rc;
    /* rows (id_query, id_other) in the deterministic order; *total = true count, rows beyond capacity dropped */
    return col_traverse_wide_pairs(b->pairs, capacity, /*base=*/0, b->total, b->wide, b->bounds, b->scene, b->ids1,
                                   COL_F32, n, /*q_begin=*/0, /*q_end=*/n, b->traverse_scratch, s);
}
