// sched.h — host-side dependency analysis of recorded reconstruction commands.
//
// The host walk (dav1d's decode_b order) emits commands in a legal serial order. A "chain"
// is the run of commands belonging to one coding block (EAV1_F_BLOCK_HEAD starts one); it
// executes on one CTA in order. Chains only depend on each other through reconstructed
// pixels of the CURRENT frame: intra edges (top / left / top-right / bottom-left strips),
// the luma block read by cfl_ac, and intra-block-copy sources. We track, per plane and per
// 4x4-pixel cell, the last chain that wrote it; a chain depends on the last writers of every
// cell it reads or rewrites. Decode order is already a topological order of that graph, so
// the device kernel hands chains out in decode order (a ticket counter) and a CTA only has
// to wait for the listed producers — this reproduces the superblock-diagonal wavefront of
// intra frames without encoding it explicitly.
#pragma once
#include <stdint.h>
#include <algorithm>
#include <vector>
#include "../../include/easyav1_cuda.h"

namespace eav1 {

struct Chain { uint32_t first, count; };

// chains in decode order + CSR producer lists (indices into chains, all smaller than the consumer)
struct DepGraph {
    std::vector<Chain> chains;
    std::vector<uint32_t> dep_start;   // chains.size() + 1
    std::vector<uint32_t> deps;
};

// Two-phase plan of a frame. Phase 1: motion compensation from reference pictures (MC with
// ref != 7, WARP) depends on nothing inside the frame, so all of it runs first as independent
// work items of at most 32x32 samples. Phase 2: every other command is a unit of the dataflow
// graph (producers = last phase-2 writers of the 4x4 cells it reads or rewrites, plus phase-2
// writers of the scratch regions it reads). Scratch regions are single-assignment (recorder).
struct UnitGraph {
    std::vector<uint32_t> mc_items;    // cmd index | tile_x << 24 | tile_y << 26
    std::vector<uint32_t> unit_cmd;    // phase-2 units in decode order: first command ...
    std::vector<uint32_t> unit_cnt;    // ... and how many consecutive commands the unit runs
    std::vector<uint32_t> dep_start;   // unit_cmd.size() + 1
    std::vector<uint32_t> deps;        // producer unit ids (smaller than the consumer)
};

// level view of the same graph (analysis / tests): level = 1 + max level of the producers
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

    void build_graph(const Eav1Cmd *cmds, size_t n, DepGraph &g) {
        g.chains.clear(); g.dep_start.clear(); g.deps.clear();
        mark_.clear();
        size_t i = 0;
        while (i < n) {
            size_t j = i + 1;
            while (j < n && !(cmds[j].flags & EAV1_F_BLOCK_HEAD)) j++;
            const uint32_t id = (uint32_t)g.chains.size();   // this chain's index
            cur_ = id + 1;
            mark_.push_back(0);
            g.dep_start.push_back((uint32_t)g.deps.size());
            out_ = &g.deps;
            // pass 1: collect the last writers of everything read or rewritten
            for (size_t k = i; k < j; k++) touch(cmds[k], false);
            // pass 2: this chain becomes the last writer of what it writes
            for (size_t k = i; k < j; k++) touch(cmds[k], true);
            Chain c; c.first = (uint32_t)i; c.count = (uint32_t)(j - i);
            g.chains.push_back(c);
            i = j;
        }
        g.dep_start.push_back((uint32_t)g.deps.size());
    }

    static bool is_ref_mc(const Eav1Cmd &c) {
        return (c.op == EAV1_OP_MC && c.u.mc.ref != 7) || c.op == EAV1_OP_WARP;
    }

    void build_units(const Eav1Cmd *cmds, size_t n, UnitGraph &g) {
        g.mc_items.clear(); g.unit_cmd.clear(); g.unit_cnt.clear(); g.dep_start.clear(); g.deps.clear();
        mark_.clear();
        prod_.clear();
        last_big_ = false;
        for (size_t i = 0; i < n; i++) {
            const Eav1Cmd &c = cmds[i];
            if (c.flags & EAV1_F_BLOCK_HEAD) prod_.clear();
            if (c.op == EAV1_OP_NOP) continue;
            if (is_ref_mc(c)) {
                bool clean = true;
                if (!(c.flags & EAV1_F_DST_ARENA)) clean = !stamped(c.plane, c.x, c.y, c.x + c.w, c.y + c.h);
                if (clean) {
                    if (c.op == EAV1_OP_WARP || c.u.mc.dx) g.mc_items.push_back((uint32_t)i);
                    else
                        for (int ty = 0; ty * 32 < c.h; ty++)
                            for (int tx = 0; tx * 32 < c.w; tx++)
                                g.mc_items.push_back((uint32_t)i | (uint32_t)tx << 24 | (uint32_t)ty << 26);
                    continue;
                }
            }
            // A command that consumes what the previous unit produced, and directly follows it in
            // the command list, joins that unit: the pair runs back to back on one work group
            // with no flag hop (prediction -> residual -> next transform block's prediction ...).
            uint32_t id = (uint32_t)g.unit_cmd.size();
            bool fuse = false;
            if (id > 0 && g.unit_cmd[id - 1] + g.unit_cnt[id - 1] == i && g.unit_cnt[id - 1] < 64 &&
                (c.w > 32 || c.h > 32) == last_big_) {
                cur_ = id;                      // probe with the previous unit's identity
                probe_hit_ = false;
                probe_ = true;
                touch(c, false);
                arena_reads(c);
                probe_ = false;
                fuse = probe_hit_;
            }
            if (fuse) {
                id -= 1;
                g.unit_cnt[id]++;
            } else {
                mark_.push_back(0);
                g.dep_start.push_back((uint32_t)g.deps.size());
                g.unit_cmd.push_back((uint32_t)i);
                g.unit_cnt.push_back(1);
                last_big_ = c.w > 32 || c.h > 32;
            }
            cur_ = id + 1;
            out_ = &g.deps;
            touch(c, false);
            arena_reads(c);
            touch(c, true);
            if (c.flags & EAV1_F_DST_ARENA)
                prod_.push_back({c.dst_off, c.dst_off + (uint32_t)c.w * c.h * 2u, id});
            if (c.op == EAV1_OP_W_MASK)
                prod_.push_back({c.u.comp.mask_off, c.u.comp.mask_off + (uint32_t)c.w * c.h, id});
        }
        g.dep_start.push_back((uint32_t)g.deps.size());
    }

    // level of every unit (analysis): 1 + max level of its producers
    static uint32_t unit_levels(const UnitGraph &g, std::vector<uint32_t> *lvl_out = nullptr) {
        std::vector<uint32_t> lvl(g.unit_cmd.size(), 0);
        uint32_t mx = 0;
        for (size_t u = 0; u < lvl.size(); u++) {
            uint32_t lv = 0;
            for (uint32_t d = g.dep_start[u]; d < g.dep_start[u + 1]; d++) lv = std::max(lv, lvl[g.deps[d]] + 1);
            lvl[u] = lv; mx = std::max(mx, lv);
        }
        if (lvl_out) lvl_out->swap(lvl);
        return g.unit_cmd.empty() ? 0 : mx + 1;
    }

    // Ticket order of the dataflow kernel: order[t] = unit that holds ticket t.
    // est = earliest possible start (longest producer path, in commands), tail = longest chain of
    // consumers behind the unit (itself included). Tickets go by w * est - tail (w >= 1), ties in decode
    // order. For a producer p of c: est[c] >= est[p] + n[p] and tail[p] >= tail[c] + n[p], so the key
    // strictly grows along every edge: producers precede consumers and a work group only ever waits for
    // smaller tickets (progress for any grid size). Among the orders with that property this one hands
    // out the frame's current wavefront, critical chains first: units on the longest path (est + tail
    // maximal) lead their level, and the many independent units (est 0, tail 1) fill in behind the heads
    // of the chains instead of in front of their second units. Larger w leans towards pure
    // earliest-start order. Two linear passes and one counting sort.
    static void ticket_order(const UnitGraph &g, uint32_t w, std::vector<uint32_t> &order, std::vector<uint32_t> &est,
                             std::vector<uint32_t> &tail, std::vector<uint32_t> &bucket) {
        const size_t nc = g.unit_cmd.size();
        est.resize(nc); order.resize(nc); tail.assign(nc, 0);
        uint32_t est_max = 0, tail_max = 0;
        for (size_t u = 0; u < nc; u++) {
            uint32_t e = 0;
            for (uint32_t d = g.dep_start[u]; d < g.dep_start[u + 1]; d++)
                e = std::max(e, est[g.deps[d]] + g.unit_cnt[g.deps[d]]);
            est[u] = e; est_max = std::max(est_max, e);
        }
        for (size_t u = nc; u-- > 0;) {
            tail[u] += g.unit_cnt[u];
            tail_max = std::max(tail_max, tail[u]);
            for (uint32_t d = g.dep_start[u]; d < g.dep_start[u + 1]; d++)
                tail[g.deps[d]] = std::max(tail[g.deps[d]], tail[u]);
        }
        bucket.assign((size_t)w * est_max + tail_max + 2, 0);
        for (size_t u = 0; u < nc; u++) bucket[w * est[u] + tail_max - tail[u] + 1]++;
        for (size_t b = 1; b < bucket.size(); b++) bucket[b] += bucket[b - 1];
        for (size_t u = 0; u < nc; u++) order[bucket[w * est[u] + tail_max - tail[u]]++] = (uint32_t)u;
    }

    static void levels(const DepGraph &g, Schedule &out) {
        const size_t nc = g.chains.size();
        std::vector<uint32_t> lvl(nc, 0);
        uint32_t max_level = 0;
        for (size_t c = 0; c < nc; c++) {
            uint32_t lv = 0;
            for (uint32_t d = g.dep_start[c]; d < g.dep_start[c + 1]; d++) lv = std::max(lv, lvl[g.deps[d]] + 1);
            lvl[c] = lv;
            max_level = std::max(max_level, lv);
        }
        out.level_start.assign(max_level + 2, 0);
        for (uint32_t l : lvl) out.level_start[l + 1]++;
        for (size_t l = 1; l < out.level_start.size(); l++) out.level_start[l] += out.level_start[l - 1];
        out.chains.resize(nc);
        std::vector<uint32_t> pos(out.level_start.begin(), out.level_start.end() - 1);
        for (size_t c = 0; c < nc; c++) out.chains[pos[lvl[c]]++] = g.chains[c];
        if (!nc) out.level_start.assign(1, 0);
    }

    void build(const Eav1Cmd *cmds, size_t n, Schedule &out) {
        DepGraph g;
        build_graph(cmds, n, g);
        levels(g, out);
    }

private:
    int cw_[3], ch_[3];
    std::vector<uint32_t> last_[3];   // per 4x4 cell: index+1 of the last chain that wrote it
    std::vector<uint32_t> mark_;      // per chain: consumer (index+1) that already listed it
    std::vector<uint32_t> *out_ = nullptr;
    uint32_t cur_ = 0;
    struct Prod { uint32_t lo, hi, unit; };
    std::vector<Prod> prod_;          // phase-2 writers of scratch regions of the current block
    bool probe_ = false, probe_hit_ = false, last_big_ = false;

    bool stamped(int p, int x0, int y0, int x1, int y1) const {
        int cx0 = std::max(x0, 0) >> 2, cy0 = std::max(y0, 0) >> 2;
        int cx1 = std::min((x1 + 3) >> 2, cw_[p]), cy1 = std::min((y1 + 3) >> 2, ch_[p]);
        for (int cy = cy0; cy < cy1; cy++)
            for (int cx = cx0; cx < cx1; cx++)
                if (last_[p][(size_t)cy * cw_[p] + cx]) return true;
        return false;
    }

    void arena_dep(uint32_t off) {
        for (const Prod &p : prod_) {
            if (off < p.lo || off >= p.hi) continue;
            if (probe_) { probe_hit_ |= p.unit + 1 == cur_; continue; }
            if (p.unit + 1 != cur_ && mark_[p.unit] != cur_) { mark_[p.unit] = cur_; out_->push_back(p.unit); }
        }
    }

    void arena_reads(const Eav1Cmd &c) {
        if (prod_.empty()) return;
        switch (c.op) {
        case EAV1_OP_AVG: case EAV1_OP_W_AVG: case EAV1_OP_MASK: case EAV1_OP_W_MASK:
            arena_dep(c.u.comp.t1); arena_dep(c.u.comp.t2);
            if (c.op == EAV1_OP_MASK && (c.flags & EAV1_F_MASK_ARENA)) arena_dep(c.u.comp.mask_off);
            break;
        case EAV1_OP_BLEND: case EAV1_OP_BLEND_V: case EAV1_OP_BLEND_H:
            arena_dep(c.u.blend.tmp_off);
            break;
        default: break;
        }
    }

    void rect(int p, int x0, int y0, int x1, int y1, bool write) {
        // [x0,x1) x [y0,y1) in pixels
        int cx0 = std::max(x0, 0) >> 2, cy0 = std::max(y0, 0) >> 2;
        int cx1 = std::min((x1 + 3) >> 2, cw_[p]), cy1 = std::min((y1 + 3) >> 2, ch_[p]);
        for (int cy = cy0; cy < cy1; cy++) {
            uint32_t *row = &last_[p][(size_t)cy * cw_[p]];
            for (int cx = cx0; cx < cx1; cx++) {
                if (write) { row[cx] = cur_; continue; }
                const uint32_t w = row[cx];
                if (probe_) { probe_hit_ |= w == cur_; continue; }
                if (w && w != cur_ && mark_[w - 1] != cur_) { mark_[w - 1] = cur_; out_->push_back(w - 1); }
            }
        }
    }

    void touch(const Eav1Cmd &c, bool write) {
        const bool dst_frame = !(c.flags & EAV1_F_DST_ARENA);
        if (dst_frame) rect(c.plane, c.x, c.y, c.x + c.w, c.y + c.h, write);
        if (write) return;
        switch (c.op) {
        case EAV1_OP_IPRED:
        case EAV1_OP_CFL: {
            const int ex = c.u.ipred.ex, ey = c.u.ipred.ey;
            const int have_left = c.u.ipred.have & 1, have_top = (c.u.ipred.have >> 1) & 1;
            if (have_top) {
                int n = c.u.ipred.n_top == 255 ? 0 : c.u.ipred.n_top;
                if (c.u.ipred.have & 16) n = (c.u.ipred.tw << 2) + c.u.ipred.n_tr;
                rect(c.plane, ex - have_left, ey - 1, ex + std::max(n, 1), ey, false);
            }
            if (have_left) {
                int n = c.u.ipred.n_left == 255 ? 0 : c.u.ipred.n_left;
                if (c.u.ipred.have & 8) n = (c.u.ipred.th << 2) + c.u.ipred.n_bl;
                rect(c.plane, ex - 1, ey - have_top, ex, ey + std::max(n, 1), false);
            }
            if (c.op == EAV1_OP_CFL) {
                // luma block feeding cfl_ac; subsampling factor is at most 2 in each direction
                rect(0, c.u.ipred.lx, c.u.ipred.ly, c.u.ipred.lx + 2 * c.w, c.u.ipred.ly + 2 * c.h, false);
            }
            break;
        }
        case EAV1_OP_MC:
            if (c.u.mc.ref == 7)
                rect(c.plane, c.u.mc.sx - 3, c.u.mc.sy - 3, c.u.mc.sx + c.w + 5, c.u.mc.sy + c.h + 5, false);
            break;
        default:
            break;
        }
    }
};

} // namespace eav1
