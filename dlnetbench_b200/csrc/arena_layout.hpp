// arena_layout.hpp — where a symmetric allocation lands inside an arena chunk (host logic only, no CUDA).
//
// Every rank runs this with the same sequence of sizes (dlnb_symm_alloc / dlnb_symm_free are
// collective), and the placement is a pure function of that sequence, so an allocation has the
// same offset in every rank's copy of the chunk — which is what lets a kernel address a peer's
// buffer as "peer window + my offset" with no table exchange. Replaces the cudaMalloc / cudaFree
// pairs of the reference's Tensor (cpp/proxy_classes.hpp:401-405,434).
#pragma once
#include <stddef.h>

#include <algorithm>
#include <vector>

namespace dlnb {

class ChunkLayout {
  public:
    explicit ChunkLayout(size_t size = 0) : size_(size) {}
    size_t size() const { return size_; }
    size_t high_water() const { return bump_; } // bytes below this offset have been handed out at some point
    size_t bytes_free_listed() const {
        size_t s = 0;
        for (const Block &b : free_) s += b.size;
        return s;
    }
    // First fit among released blocks (lowest offset first), else the untouched tail.
    // `recycled` tells the caller that the block has held data before (it must be cleared again).
    bool take(size_t need, size_t *off, bool *recycled) {
        for (size_t i = 0; i < free_.size(); ++i) {
            if (free_[i].size < need) continue;
            *off = free_[i].off;
            *recycled = true;
            if (free_[i].size == need) free_.erase(free_.begin() + (long)i);
            else { free_[i].off += need; free_[i].size -= need; }
            return true;
        }
        if (bump_ + need > size_) return false;
        *off = bump_;
        *recycled = bump_ < dirty_below_;
        bump_ += need;
        return true;
    }
    // Give a block back: merged with its neighbours; a block that ends at the high-water mark
    // lowers the mark instead (so a last-in-first-out caller never fragments the chunk).
    void give_back(size_t off, size_t size) {
        dirty_below_ = std::max(dirty_below_, off + size);
        Block nb{off, size};
        auto it = std::lower_bound(free_.begin(), free_.end(), nb, [](const Block &a, const Block &b) { return a.off < b.off; });
        it = free_.insert(it, nb);
        if (it + 1 != free_.end() && it->off + it->size == (it + 1)->off) { // merge with the right neighbour
            it->size += (it + 1)->size;
            free_.erase(it + 1);
        }
        if (it != free_.begin() && (it - 1)->off + (it - 1)->size == it->off) { // and with the left one
            (it - 1)->size += it->size;
            it = free_.erase(it) - 1;
        }
        if (it->off + it->size == bump_) { // trailing block: back into the untouched-tail bookkeeping
            bump_ = it->off;
            free_.erase(it);
        }
    }

  private:
    struct Block {
        size_t off, size;
    };
    size_t size_ = 0, bump_ = 0, dirty_below_ = 0;
    std::vector<Block> free_; // sorted by offset, no two adjacent
};

} // namespace dlnb
