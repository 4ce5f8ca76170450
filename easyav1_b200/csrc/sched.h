// sched.h — host-side dependency scheduling of recorded reconstruction commands.
//
// The host walk (dav1d's decode_b order) emits commands in a legal serial order. A "chain"
// is the run of commands belonging to one coding block (EAV1_F_BLOCK_HEAD starts one); it
// executes on one CTA in order. Chains only depend on each other through reconstructed
// pixels of the CURRENT frame: intra edges (top / left / top-right / bottom-left strips),
// the luma block read by cfl_ac, and intra-block-copy sources. We track, per plane and per
// 4x4-pixel cell, the level of the last chain that wrote it; a chain's level is one more
// than the highest level among the cells it reads or rewrites. Chains of equal level are
// independent and run as one kernel launch; this reproduces the superblock-diagonal
// wavefront of intra frames without encoding it explicitly.
#pragma once
#include <stdint.h>
#include <algorithm>
#include <vector>
#include "../../include/easyav1_cuda.h"

namespace eav1 {

struct Chain { uint32_t first, count; };

struct Schedule {
    std::vector<Chain> chains;          // sorted by level
    std::vector<uint32_t> level_start;  // level l covers chains[level_start[l] .. level_start[l+1])
};

class Scheduler {
public:
    void reset(const int pw[3], const int ph[3]) {
        for (int p = 0; p < 3; p++) {
            cw_[p] = (pw[p] + 3) >> 2;
            ch_[p] = (ph[p] + 3) >> 2;
            last_[p].assign((size_t)cw_[p] * ch_[p], 0);
        }
    }

    void build(const Eav1Cmd *cmds, size_t n, Schedule &out) {
        tmp_.clear();
        lvl_.clear();
        size_t i = 0;
        uint32_t max_level = 0;
        while (i < n) {
            size_t j = i + 1;
            while (j < n && !(cmds[j].flags & EAV1_F_BLOCK_HEAD)) j++;
            // pass 1: level = 1 + max(last writer of everything read or written)
            uint32_t lv = 0;
            for (size_t k = i; k < j; k++) lv = std::max(lv, touch(cmds[k], 0, false));
            lv += 1;
            // pass 2: stamp written cells
            for (size_t k = i; k < j; k++) touch(cmds[k], lv, true);
            Chain c; c.first = (uint32_t)i; c.count = (uint32_t)(j - i);
            tmp_.push_back(c);
            lvl_.push_back(lv);
            max_level = std::max(max_level, lv);
            i = j;
        }
        // counting sort by level
        out.level_start.assign(max_level + 2, 0);
        for (uint32_t l : lvl_) out.level_start[l + 1]++;
        for (size_t l = 1; l < out.level_start.size(); l++) out.level_start[l] += out.level_start[l - 1];
        out.chains.resize(tmp_.size());
        std::vector<uint32_t> pos(out.level_start.begin(), out.level_start.end() - 1);
        for (size_t c = 0; c < tmp_.size(); c++) out.chains[pos[lvl_[c]]++] = tmp_[c];
        // drop the empty level 0 slot
        out.level_start.erase(out.level_start.begin());
    }

private:
    int cw_[3], ch_[3];
    std::vector<uint32_t> last_[3];
    std::vector<Chain> tmp_;
    std::vector<uint32_t> lvl_;

    uint32_t rect(int p, int x0, int y0, int x1, int y1, uint32_t stamp, bool write) {
        // [x0,x1) x [y0,y1) in pixels
        int cx0 = std::max(x0, 0) >> 2, cy0 = std::max(y0, 0) >> 2;
        int cx1 = std::min((x1 + 3) >> 2, cw_[p]), cy1 = std::min((y1 + 3) >> 2, ch_[p]);
        uint32_t m = 0;
        for (int cy = cy0; cy < cy1; cy++) {
            uint32_t *row = &last_[p][(size_t)cy * cw_[p]];
            for (int cx = cx0; cx < cx1; cx++) {
                if (write) row[cx] = stamp;
                else m = std::max(m, row[cx]);
            }
        }
        return m;
    }

    // write == false: returns max level of cells read or written; write == true: stamps writes
    uint32_t touch(const Eav1Cmd &c, uint32_t stamp, bool write) {
        uint32_t m = 0;
        const bool dst_frame = !(c.flags & EAV1_F_DST_ARENA);
        if (dst_frame) m = std::max(m, rect(c.plane, c.x, c.y, c.x + c.w, c.y + c.h, stamp, write));
        if (write) return m;
        switch (c.op) {
        case EAV1_OP_IPRED:
        case EAV1_OP_CFL: {
            const int ex = c.u.ipred.ex, ey = c.u.ipred.ey;
            const int have_left = c.u.ipred.have & 1, have_top = (c.u.ipred.have >> 1) & 1;
            if (have_top) {
                int n = c.u.ipred.n_top == 255 ? 0 : c.u.ipred.n_top;
                if (c.u.ipred.have & 16) n = (c.u.ipred.tw << 2) + c.u.ipred.n_tr;
                m = std::max(m, rect(c.plane, ex - have_left, ey - 1, ex + std::max(n, 1), ey, 0, false));
            }
            if (have_left) {
                int n = c.u.ipred.n_left == 255 ? 0 : c.u.ipred.n_left;
                if (c.u.ipred.have & 8) n = (c.u.ipred.th << 2) + c.u.ipred.n_bl;
                m = std::max(m, rect(c.plane, ex - 1, ey - have_top, ex, ey + std::max(n, 1), 0, false));
            }
            if (c.op == EAV1_OP_CFL) {
                // luma block feeding cfl_ac; subsampling factor is at most 2 in each direction
                m = std::max(m, rect(0, c.u.ipred.lx, c.u.ipred.ly, c.u.ipred.lx + 2 * c.w, c.u.ipred.ly + 2 * c.h, 0, false));
            }
            break;
        }
        case EAV1_OP_MC:
            if (c.u.mc.ref == 7)
                m = std::max(m, rect(c.plane, c.u.mc.sx - 3, c.u.mc.sy - 3, c.u.mc.sx + c.w + 5, c.u.mc.sy + c.h + 5, 0, false));
            break;
        default:
            break;
        }
        return m;
    }
};

} // namespace eav1
