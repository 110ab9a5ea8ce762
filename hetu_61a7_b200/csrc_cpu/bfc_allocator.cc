// Best-fit-with-coalescing sub-allocator over one contiguous arena (host-side bookkeeping only: it hands
// out OFFSETS, the caller adds them to whatever base pointer the arena has - device memory, the CUDA-IPC
// symmetric heap, pinned host memory).
//
// reference: src/memory_pool/{allocator.h,BFC_allocator.h} - TF-style BFC allocator headers that are not
// compiled into any target upstream.  Here it backs the symmetric-heap arena
// (parallel/symm.py: every rank performs the same allocation sequence, so offsets are identical on all
// GPUs) and the executor's memory planner (memory_pool.py simulates a step's buffer lifetime with it).
#include <cstdint>
#include <map>
#include <mutex>
#include <set>

#define API extern "C" __attribute__((visibility("default")))

namespace {
struct BFC {
    uint64_t capacity, align;
    std::map<uint64_t, uint64_t> free_by_off;                 // offset -> size
    std::set<std::pair<uint64_t, uint64_t>> free_by_size;     // (size, offset)
    std::map<uint64_t, uint64_t> used;                        // offset -> size
    uint64_t in_use = 0, peak = 0, n_alloc = 0;
    std::mutex mu;

    BFC(uint64_t cap, uint64_t al) : capacity(cap), align(al ? al : 256) {
        free_by_off[0] = cap;
        free_by_size.insert({cap, 0});
    }
    void drop_free(uint64_t off, uint64_t size) {
        free_by_off.erase(off);
        free_by_size.erase({size, off});
    }
    void add_free(uint64_t off, uint64_t size) {
        free_by_off[off] = size;
        free_by_size.insert({size, off});
    }
    int64_t alloc(uint64_t bytes) {
        std::lock_guard<std::mutex> l(mu);
        uint64_t need = (bytes + align - 1) / align * align;
        if (need == 0) need = align;
        auto it = free_by_size.lower_bound({need, 0});        // smallest block that fits
        if (it == free_by_size.end()) return -1;
        uint64_t size = it->first, off = it->second;
        drop_free(off, size);
        if (size > need) add_free(off + need, size - need);   // split, keep the tail free
        used[off] = need;
        in_use += need;
        if (in_use > peak) peak = in_use;
        ++n_alloc;
        return (int64_t)off;
    }
    int free_block(uint64_t off) {
        std::lock_guard<std::mutex> l(mu);
        auto u = used.find(off);
        if (u == used.end()) return -1;
        uint64_t size = u->second;
        used.erase(u);
        in_use -= size;
        auto next = free_by_off.find(off + size);             // coalesce with the right neighbour
        if (next != free_by_off.end()) {
            uint64_t nsize = next->second;
            drop_free(off + size, nsize);
            size += nsize;
        }
        auto right = free_by_off.lower_bound(off);            // coalesce with the left neighbour
        if (right != free_by_off.begin()) {
            auto left = std::prev(right);
            if (left->first + left->second == off) {
                uint64_t loff = left->first, lsize = left->second;
                drop_free(loff, lsize);
                off = loff;
                size += lsize;
            }
        }
        add_free(off, size);
        return 0;
    }
};
}  // namespace

API void* hetu_bfc_create(uint64_t capacity, uint64_t alignment) { return new BFC(capacity, alignment); }
API void hetu_bfc_destroy(void* h) { delete static_cast<BFC*>(h); }
API int64_t hetu_bfc_alloc(void* h, uint64_t bytes) { return static_cast<BFC*>(h)->alloc(bytes); }
API int hetu_bfc_free(void* h, uint64_t offset) { return static_cast<BFC*>(h)->free_block(offset); }
// out[0] bytes in use, out[1] peak bytes in use, out[2] live blocks, out[3] free fragments,
// out[4] largest free block, out[5] allocations served
API void hetu_bfc_stats(void* h, uint64_t* out) {
    BFC* b = static_cast<BFC*>(h);
    std::lock_guard<std::mutex> l(b->mu);
    out[0] = b->in_use;
    out[1] = b->peak;
    out[2] = b->used.size();
    out[3] = b->free_by_off.size();
    out[4] = b->free_by_size.empty() ? 0 : b->free_by_size.rbegin()->first;
    out[5] = b->n_alloc;
}
