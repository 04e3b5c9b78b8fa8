// K4: fused MM energy + analytic dE/dxyz, K5: its VJP w.r.t. the MM parameters.  See include/grappa_b200.h.
//
// Data layout in HBM: xyz / gradient are (N, C, 3) fp32, atom-major, conformation-minor (the reference's layout,
// internal_coordinates.py:31), so the coordinates of one atom for 32 consecutive conformations are 384 contiguous
// bytes.  One CTA evaluates (atom block, chunk of <= 32 conformations) with LANE = CONFORMATION:
//   phase 0  the block's coordinate rows are staged into shared memory by the TMA engine (one cp.async.bulk per atom
//            row, completion on an mbarrier) while the threads copy the block's tuple PROGRAM (below) and the
//            molecule's MM parameters into shared memory and clear the gradient accumulators.
//   phase 1  each of the W warps walks its own part of the block's program.  Every record is warp-uniform (all lanes
//            decode the same tuple from a broadcast shared-memory read, no divergence); lane c evaluates it for
//            conformation c: positions are conflict-free reads (stride 3 floats across lanes), the analytic gradient
//            vectors are added into the warp's PRIVATE accumulator rows [slot][conformation][xyz] (again stride 3,
//            conflict-free, no atomics, program order = summation order).
//   phase 2  the W private accumulators are summed in warp order into the coordinate buffer, which then leaves as one
//            cp.async.bulk store per atom row; per-lane energy partials are summed in warp order.
// Everything is a fixed-order sum -> forces and energies are bit-reproducible.
// Algorithmic HBM bytes per (molecule, conformation) with A atoms: 12A (xyz) + 12A (gradient) + 20 (energies).
#include "common.cuh"
#include "geometry.cuh"
#include "energy_plan.h"
#include <vector>
#include <algorithm>
#include <mutex>

namespace grappa {

// ------------------------------------------------------------------------------------------------ host plan ----
//
// Block row (12 ints):  0 a0 | 1 n_atoms | 2 molecule | 3 prog0 | 4 n_words | 5 n_slots | 6 blocks of this molecule |
//                       7 ordinal of this block in its molecule | 8 local (every atom of every record is inside the
//                       block) | 9..11 unused.  Columns 0 / 2 / 3 are absolute (atom row, molecule, word offset); every
//                       word of the program is relative to the block or to the molecule, so per-molecule plans
//                       concatenate into a batch plan by adding three offsets per block row (csrc/collate.cu, mode 1).
// Program of a block:   [off_0 .. off_W]  word offsets (relative to prog0) of the W parts,
//                       [pull_ptr[al]]    n_atoms + 1 offsets into the pull list (relative to its start),
//                       [pull[..]]        per atom, in part order: part << 16 | 96 * slot  = the accumulator rows whose
//                                         sum is the atom's gradient (phase 2 of the kernel),
//                       then per part [n_proper, n_improper, n_angle, n_bond] followed by its records in that order:
//                         torsion  [hdr, w0, w1, w2, w3]     hdr = t_local << 1 | owner
//                         angle    [hdr, w0, w1, w2]
//                         bond     [hdr, w0, w1]
//                       t_local = tuple row minus the first row of the block's molecule at that level (mol_tuple_off);
//                       owner = this block adds the tuple's energy (the block of its first atom).
//                       Atom word w: local block  -> (384 * local atom) | (384 * accumulator slot) << 16   (byte offsets)
//                                    other blocks -> atom row minus a0 (signed; outside [0, n_atoms) = halo atom read
//                                                    from global memory, no gradient accumulated here); slot = atom.
// Parts: the block's tuples are sorted by the mean index of their atoms and cut into W runs of equal estimated cost,
// so a part touches a compact range of the molecule; its accumulator has one slot per atom it touches (local blocks),
// which is what keeps the per-warp shared-memory footprint small enough for many resident warps.
constexpr int kBlkStride = 12;
constexpr int kMaxParts = 16;
constexpr int kRow = 96;          // floats per (atom, 32 conformations) row

struct LevelIdx {
    const int64_t* idx;
    int64_t T;
    int n_seq;
};

struct PlanItem {
    int level;       // 0 bond, 1 angle, 2 proper, 3 improper
    int64_t t;
    int64_t key;     // 12 x mean atom index
};

}  // namespace grappa

using namespace grappa;

int grappa::build_energy_plan(int64_t n_atoms, int64_t n_mol, const int64_t* mol_atom_off,
                              const int64_t* idx2, int64_t T2, const int64_t* idx3, int64_t T3,
                              const int64_t* idx4, int64_t T4, const int64_t* idx4i, int64_t T4i,
                              int block_atoms, int n_parts, bool fill, EnergyPlanOut& out) {
    int64_t* sizes_out = out.sizes;
    out.tuple_mol.assign((size_t)(T2 + T3 + T4 + T4i), 0);
    int32_t* tuple_mol = out.tuple_mol.data();
    GRAPPA_CHECK_ARG(n_atoms >= 0 && n_mol >= 0, "bad sizes");
    GRAPPA_CHECK_ARG(block_atoms > 0 && block_atoms <= 128, "block_atoms must be in [1, 128]");
    GRAPPA_CHECK_ARG(n_parts >= 1 && n_parts <= kMaxParts, "n_parts must be in [1, 16]");
    GRAPPA_CHECK_ARG(mol_atom_off != nullptr, "null pointer");
    GRAPPA_CHECK_ARG(mol_atom_off[0] == 0 && mol_atom_off[n_mol] == n_atoms, "mol_atom_off must span [0, n_atoms]");
    LevelIdx lv[4] = {{idx2, T2, 2}, {idx3, T3, 3}, {idx4, T4, 4}, {idx4i, T4i, 4}};
    const int W = n_parts;
    const int kCost[4] = {6, 14, 18, 18};              // relative cost of a bond / angle / torsion record (measured SASS instructions / 10)
    const int kOrder[4] = {3, 2, 0, 1};               // position of a level inside a part: propers, impropers, angles, bonds
    const int kWords[4] = {3, 4, 5, 5};

    // atom -> block, block table
    std::vector<int32_t> atom_blk((size_t)n_atoms), atom_mol((size_t)n_atoms);
    struct Blk { int32_t a0, na, mol, nb, j; };
    std::vector<Blk> blocks;
    for (int64_t m = 0; m < n_mol; ++m) {
        int64_t a0 = mol_atom_off[m], A = mol_atom_off[m + 1] - a0;
        GRAPPA_CHECK_ARG(A >= 0, "mol_atom_off must be non-decreasing");
        int64_t nb = std::max<int64_t>(1, ceil_div(A, block_atoms));
        int64_t per = ceil_div(A, nb);
        for (int64_t j = 0; j < nb; ++j) {
            int64_t s = a0 + j * per, e = std::min(a0 + A, s + per);
            if (e < s) e = s;
            for (int64_t a = s; a < e; ++a) { atom_blk[a] = (int32_t)blocks.size(); atom_mol[a] = (int32_t)m; }
            blocks.push_back({(int32_t)s, (int32_t)(e - s), (int32_t)m, (int32_t)nb, (int32_t)j});
        }
    }
    const int64_t nB = (int64_t)blocks.size();

    // tuples: bounds, molecule of each tuple (= molecule of its first atom), per-molecule row offsets per level.
    // Rows of a level must be grouped by molecule in molecule order (what dgl.batch / graph.batch produce).
    std::vector<int64_t> toff((size_t)4 * (n_mol + 1), 0);
    int64_t maxT[4] = {0, 0, 0, 0};
    int64_t tuple_base = 0;
    for (int l = 0; l < 4; ++l) {
        int64_t* to = toff.data() + (size_t)l * (n_mol + 1);
        int32_t prev = 0;
        for (int64_t t = 0; t < lv[l].T; ++t) {
            const int64_t* a = lv[l].idx + t * lv[l].n_seq;
            for (int i = 0; i < lv[l].n_seq; ++i)
                if (a[i] < 0 || a[i] >= n_atoms)
                    return fail("grappa_energy_plan_build: %s atom index %lld out of range [0,%lld)", "", (long long)a[i], (long long)n_atoms);
            const int32_t m = atom_mol[a[0]];
            if (m < prev) return fail("grappa_energy_plan_build: %s tuple rows must be grouped by molecule in molecule order (row %lld)", "", (long long)t);
            prev = m;
            to[m + 1]++;
            tuple_mol[tuple_base + t] = m;
        }
        for (int64_t m = 0; m < n_mol; ++m) { maxT[l] = std::max(maxT[l], to[m + 1]); to[m + 1] += to[m]; }
        tuple_base += lv[l].T;
    }

    // per block: the tuples touching it
    std::vector<std::vector<PlanItem>> items((size_t)nB);
    for (int l = 0; l < 4; ++l) {
        for (int64_t t = 0; t < lv[l].T; ++t) {
            const int64_t* a = lv[l].idx + t * lv[l].n_seq;
            int32_t tb[4]; int nt = 0;
            int64_t sum = 0;
            for (int i = 0; i < lv[l].n_seq; ++i) {
                int32_t b = atom_blk[a[i]];
                bool seen = false;
                for (int q = 0; q < nt; ++q) seen |= (tb[q] == b);
                if (!seen) tb[nt++] = b;
                sum += a[i];
            }
            for (int q = 0; q < nt; ++q) items[(size_t)tb[q]].push_back({l, t, sum * 12 / lv[l].n_seq});
        }
    }

    // per block: locality, parts (cut of the key-sorted list into W runs of equal cost), slots, program size
    int64_t max_words = 0, max_atoms = 0, max_slots = 1, any_multi = 0, any_nonlocal = 0;
    std::vector<int64_t> word0((size_t)nB + 1, 0);
    std::vector<uint8_t> local((size_t)nB, 1);
    std::vector<std::vector<int32_t>> part_of((size_t)nB);     // part of every item (after sorting)
    std::vector<int32_t> n_slots((size_t)nB, 0), n_pull((size_t)nB, 0);
    for (int64_t b = 0; b < nB; ++b) {
        std::vector<PlanItem>& v = items[(size_t)b];
        std::stable_sort(v.begin(), v.end(), [](const PlanItem& x, const PlanItem& y) { return x.key < y.key; });
        int64_t total = 0, w = (W + 1) + (blocks[b].na + 1) + 4 * W;
        for (const PlanItem& it : v) {
            total += kCost[it.level];
            w += kWords[it.level];
            const int64_t* a = lv[it.level].idx + it.t * lv[it.level].n_seq;
            for (int i = 0; i < lv[it.level].n_seq; ++i) if (atom_blk[a[i]] != b) local[b] = 0;
        }
        std::vector<int32_t>& po = part_of[(size_t)b];
        po.resize(v.size());
        int64_t run = 0;
        for (size_t i = 0; i < v.size(); ++i) {
            const int c = kCost[v[i].level];
            po[i] = (int32_t)std::min<int64_t>(W - 1, (2 * run + c) * W / (2 * std::max<int64_t>(total, 1)));
            run += c;
        }
        if (local[b]) {
            int32_t ns = 0, np_ = 0;
            std::vector<uint8_t> touched((size_t)W * std::max(1, (int)blocks[b].na), 0);
            std::vector<int32_t> cnt((size_t)W, 0);
            for (size_t i = 0; i < v.size(); ++i) {
                const int64_t* a = lv[v[i].level].idx + v[i].t * lv[v[i].level].n_seq;
                for (int q = 0; q < lv[v[i].level].n_seq; ++q) {
                    uint8_t& f = touched[(size_t)po[i] * blocks[b].na + (a[q] - blocks[b].a0)];
                    if (!f) { f = 1; cnt[po[i]]++; }
                }
            }
            for (int p = 0; p < W; ++p) { ns = std::max(ns, cnt[p]); np_ += cnt[p]; }
            n_slots[b] = std::max(ns, 1);
            n_pull[b] = np_;
        } else {
            n_slots[b] = std::max(1, (int)blocks[b].na);
            n_pull[b] = W * blocks[b].na;
        }
        w += n_pull[b];
        w = (w + 63) / 64 * 64;     // coarse granularity: similar molecules get equal program sizes, so batches drawn from
                                    // them keep one buffer layout (= one captured step graph)
        word0[b + 1] = word0[b] + w;
        max_words = std::max(max_words, w);
        max_atoms = std::max<int64_t>(max_atoms, blocks[b].na);
        max_slots = std::max<int64_t>(max_slots, n_slots[b]);
        any_multi |= (blocks[b].nb > 1);
        any_nonlocal |= !local[b];
    }
    const int64_t n_words = word0[nB];
    GRAPPA_CHECK_ARG(n_words < (1LL << 31), "program too large");
    sizes_out[0] = nB; sizes_out[1] = n_words; sizes_out[2] = max_words; sizes_out[3] = max_slots;
    sizes_out[4] = max_atoms; sizes_out[5] = any_multi; sizes_out[6] = W;
    for (int l = 0; l < 4; ++l) sizes_out[7 + l] = maxT[l];
    sizes_out[11] = any_nonlocal;
    if (!fill) return 0;            // size query only
    out.blk.assign((size_t)nB * kBlkStride, 0);
    out.prog.assign((size_t)n_words, 0);
    out.toff.assign(toff.size(), 0);
    int32_t* blk = out.blk.data();
    int32_t* prog = out.prog.data();
    int32_t* mol_tuple_off = out.toff.data();

    for (size_t i = 0; i < toff.size(); ++i) mol_tuple_off[i] = (int32_t)toff[i];
    std::vector<int32_t> slot_of;
    for (int64_t b = 0; b < nB; ++b) {
        int32_t* o = blk + b * kBlkStride;
        const Blk& B = blocks[b];
        const std::vector<PlanItem>& v = items[(size_t)b];
        const std::vector<int32_t>& po = part_of[(size_t)b];
        o[0] = B.a0; o[1] = B.na; o[2] = B.mol; o[3] = (int32_t)word0[b]; o[4] = (int32_t)(word0[b + 1] - word0[b]);
        o[5] = n_slots[b]; o[6] = B.nb; o[7] = B.j; o[8] = local[b]; o[9] = o[10] = o[11] = 0;
        int32_t* pg = prog + word0[b];
        int32_t* pull_ptr = pg + (W + 1);
        int32_t* pull = pull_ptr + (B.na + 1);
        std::vector<int32_t> map((size_t)W * std::max(1, (int)B.na), -1);   // [w][al] -> 96 * slot
        if (!local[b]) for (int64_t i = 0; i < (int64_t)W * B.na; ++i) map[i] = (int32_t)((i % B.na) * kRow);
        int64_t pos = (W + 1) + (B.na + 1) + n_pull[b];
        size_t i0 = 0;
        for (int w = 0; w < W; ++w) {
            size_t i1 = i0;
            while (i1 < v.size() && po[i1] == w) ++i1;
            pg[w] = (int32_t)pos;
            int32_t* head = pg + pos;
            pos += 4;
            head[0] = head[1] = head[2] = head[3] = 0;
            int32_t next_slot = 0;
            for (int ord = 0; ord < 4; ++ord) {
                for (size_t i = i0; i < i1; ++i) {
                    const PlanItem& it = v[i];
                    if (kOrder[it.level] != ord) continue;
                    head[ord]++;
                    const int ns = lv[it.level].n_seq;
                    const int64_t* a = lv[it.level].idx + it.t * ns;
                    const int64_t tl = it.t - toff[(size_t)it.level * (n_mol + 1) + B.mol];
                    const int owner = atom_blk[a[0]] == b;
                    pg[pos++] = (int32_t)(tl * 2 + owner);
                    for (int q = 0; q < ns; ++q) {
                        const int64_t al = a[q] - B.a0;
                        if (local[b]) {
                            int32_t& m = map[(size_t)w * B.na + al];
                            if (m < 0) m = (next_slot++) * kRow;
                            pg[pos++] = (int32_t)((al * kRow * 4) | ((int64_t)m * 4 << 16));
                        } else {
                            pg[pos++] = (int32_t)al;
                        }
                    }
                }
            }
            i0 = i1;
        }
        int32_t np_ = 0;
        for (int64_t al = 0; al < B.na; ++al) {
            pull_ptr[al] = np_;
            for (int w = 0; w < W; ++w)
                if (map[(size_t)w * B.na + al] >= 0) pull[np_++] = (w << 16) | map[(size_t)w * B.na + al];
        }
        pull_ptr[B.na] = np_;
        pg[W] = (int32_t)pos;
        while (pos < word0[b + 1] - word0[b]) pg[pos++] = 0;
    }
    return 0;
}

extern "C" int grappa_energy_plan_build(int64_t n_atoms, int64_t n_mol, const int64_t* mol_atom_off,
                                        const int64_t* idx2, int64_t T2, const int64_t* idx3, int64_t T3,
                                        const int64_t* idx4, int64_t T4, const int64_t* idx4i, int64_t T4i,
                                        int block_atoms, int n_parts, int64_t* sizes_out,
                                        int32_t* blk, int32_t* prog, int32_t* mol_tuple_off, int32_t* tuple_mol) {
    GRAPPA_CHECK_ARG(sizes_out != nullptr, "null pointer");
    GRAPPA_CHECK_ARG(blk == nullptr || (prog && mol_tuple_off), "null output array");
    EnergyPlanOut out;
    if (int rc = build_energy_plan(n_atoms, n_mol, mol_atom_off, idx2, T2, idx3, T3, idx4, T4, idx4i, T4i, block_atoms, n_parts,
                                   blk != nullptr, out)) return rc;
    for (int i = 0; i < 12; ++i) sizes_out[i] = out.sizes[i];
    if (blk == nullptr) return 0;   // size query only
    std::copy(out.blk.begin(), out.blk.end(), blk);
    std::copy(out.prog.begin(), out.prog.end(), prog);
    std::copy(out.toff.begin(), out.toff.end(), mol_tuple_off);
    if (tuple_mol) std::copy(out.tuple_mol.begin(), out.tuple_mol.end(), tuple_mol);
    return 0;
}

// --------------------------------------------------------------------------------------------- device code ----

namespace grappa {

struct EnergyParams {
    const float* xyz;
    const int32_t* blk;
    const int32_t* prog;
    const int32_t* mol_toff;     // [4][n_mol + 1]
    const float* k[4];
    const float* eq[2];
    int P4, P4i;
    float* energy;
    float* term_energy;
    float* gradient;
    float* x[4];
    float* e[4];
    float* partial;
    int64_t n_mol, C;
    int term_mask;
    int max_block_atoms, max_slots, max_words, prm_floats;
    int cs;                      // conformations per chunk (<= 32)
    int use_tma;
};

__device__ __forceinline__ V3 load3(const float* __restrict__ p) { return v3(__ldg(p), __ldg(p + 1), __ldg(p + 2)); }

constexpr int kMaxPeriodicity = 6;

__device__ __forceinline__ uint32_t e_smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void e_mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(e_smem_u32(bar)), "r"(count));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void e_mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(e_smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void e_mbar_wait(uint64_t* bar, uint32_t parity) {
    uint32_t done;
    do {
        asm volatile(
            "{\n\t"
            ".reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t"
            "}\n" : "=r"(done) : "r"(e_smem_u32(bar)), "r"(parity) : "memory");
    } while (!done);
}
// TMA engine, non-tensor form: one contiguous row global -> shared (completion counted in bytes on the mbarrier) ...
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(e_smem_u32(dst_smem)), "l"(src), "r"(bytes), "r"(e_smem_u32(bar)) : "memory");
}
// ... and shared -> global (bulk async-group completion).
__device__ __forceinline__ void bulk_s2g(void* dst, const void* src_smem, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(e_smem_u32(src_smem)), "r"(bytes) : "memory");
}

// torsion energy sum_n k_n cos(n phi) and dE/dphi = -sum_n n k_n sin(n phi) from (cos phi, sin phi)
template <int P>
__device__ __forceinline__ void torsion_terms(const float* __restrict__ k, float c, float s, float& E, float& dEdphi) {
    float cn = c, sn = s;
    E = 0.f; dEdphi = 0.f;
#pragma unroll
    for (int n = 1; n <= P; ++n) {
        const float kn = k[n - 1];
        E = fmaf(kn, cn, E);
        dEdphi = fmaf(-(float)n * kn, sn, dEdphi);
        if (n < P) {
            const float c2 = cn * c - sn * s;
            sn = sn * c + cn * s;
            cn = c2;
        }
    }
}

// Everything a record evaluation needs besides the record itself.
struct LaneCtx {
    const char* __restrict__ xl;        // shared coordinates + 12 * lane bytes
    char* __restrict__ accw;            // this warp's accumulator + 12 * lane bytes (never aliases xl / the program)
    const float* __restrict__ halo;     // xyz + ((a0 * C + c0 + lane) * 3: positions of atoms outside the block
    int64_t C3, C, col;
    int na;
    bool blk_local;
};

// Atom word of a program record -> position for this lane's conformation, and the accumulator row of the atom
// (nullptr: halo atom, no gradient here).
template <bool LOCAL>
__device__ __forceinline__ V3 rec_pos(int w, const LaneCtx& c, float*& acc_row) {
    if (LOCAL || c.blk_local) {
        const float* q = reinterpret_cast<const float*>(c.xl + (w & 0xffff));
        acc_row = reinterpret_cast<float*>(c.accw + ((unsigned)w >> 16));
        return v3(q[0], q[1], q[2]);
    }
    if ((unsigned)w < (unsigned)c.na) {
        const float* q = reinterpret_cast<const float*>(c.xl + w * (kRow * 4));
        acc_row = reinterpret_cast<float*>(c.accw + w * (kRow * 4));
        return v3(q[0], q[1], q[2]);
    }
    acc_row = nullptr;
    const float* q = c.halo + (int64_t)w * c.C3;
    return v3(__ldg(q), __ldg(q + 1), __ldg(q + 2));
}

template <bool LOCAL>
__device__ __forceinline__ void acc3(float* row, float s, V3 g) {
    if (LOCAL || row != nullptr) { row[0] = fmaf(s, g.x, row[0]); row[1] = fmaf(s, g.y, row[1]); row[2] = fmaf(s, g.z, row[2]); }
}

template <int P, bool LOCAL, bool MAT>
__device__ __forceinline__ float torsion_loop(int n, const int* __restrict__ pw, const LaneCtx& c, const float* __restrict__ kp,
                                              int toff, float* __restrict__ xo, float* __restrict__ eo) {
    float esum = 0.f;
    for (int i = 0; i < n; ++i, pw += 5) {
        const int hdr = pw[0];
        float *r0, *r1, *r2, *r3;
        const V3 x0 = rec_pos<LOCAL>(pw[1], c, r0);
        const V3 x1 = rec_pos<LOCAL>(pw[2], c, r1);
        const V3 x2 = rec_pos<LOCAL>(pw[3], c, r2);
        const V3 x3 = rec_pos<LOCAL>(pw[4], c, r3);
        float cphi, sphi; V3 gg[4];
        dihedral_geom(x0, x1, x2, x3, cphi, sphi, gg);
        const int tl = hdr >> 1;
        float E, dE;
        torsion_terms<P>(kp + (int64_t)tl * P, cphi, sphi, E, dE);
        acc3<LOCAL>(r0, dE, gg[0]);
        acc3<LOCAL>(r1, dE, gg[1]);
        acc3<LOCAL>(r2, dE, gg[2]);
        acc3<LOCAL>(r3, dE, gg[3]);
        if (LOCAL || (hdr & 1)) {
            esum += E;
            if (MAT) {
                const int64_t row = (int64_t)(toff + tl) * c.C + c.col;
                if (xo) xo[row] = atan2f(sphi, cphi);
                if (eo) eo[row] = E;
            }
        }
    }
    return esum;
}

template <bool LOCAL, bool MAT>
__device__ __forceinline__ float torsion_dispatch(int P, int n, const int* pw, const LaneCtx& c, const float* kp, int toff,
                                                  float* xo, float* eo) {
    switch (P) {
        case 1: return torsion_loop<1, LOCAL, MAT>(n, pw, c, kp, toff, xo, eo);
        case 2: return torsion_loop<2, LOCAL, MAT>(n, pw, c, kp, toff, xo, eo);
        case 3: return torsion_loop<3, LOCAL, MAT>(n, pw, c, kp, toff, xo, eo);
        case 4: return torsion_loop<4, LOCAL, MAT>(n, pw, c, kp, toff, xo, eo);
        case 5: return torsion_loop<5, LOCAL, MAT>(n, pw, c, kp, toff, xo, eo);
        case 6: return torsion_loop<6, LOCAL, MAT>(n, pw, c, kp, toff, xo, eo);
        default: return 0.f;
    }
}

constexpr int kMaxEnergyThreads = 32 * kMaxParts;

template <bool LOCAL, bool MAT>
__global__ void __launch_bounds__(kMaxEnergyThreads) energy_fwd_kernel(const EnergyParams p) {
    extern __shared__ __align__(128) float smem[];
    const int W = blockDim.x >> 5, NT = blockDim.x;
    const int A96 = p.max_block_atoms * kRow, S96 = p.max_slots * kRow;
    float* xs = smem;                                  // [block atom][32 conformations][3]   (TMA destination / source)
    float* acc = xs + A96;                             // [W][slot][32][3]
    float* ered = acc + W * S96;                       // [W][4][32]
    float* prm = ered + W * 128;                       // LOCAL: k2 | eq2 | k3 | eq3 | k4 | k4i of the block's molecule
    int* prog = reinterpret_cast<int*>(prm + p.prm_floats);
    uint64_t* mbar = reinterpret_cast<uint64_t*>(prog + p.max_words);

    const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int32_t* bi = p.blk + (int64_t)b * kBlkStride;
    const int a0 = bi[0], na = bi[1], mol = bi[2], prog0 = bi[3], n_words = bi[4], n_slots = bi[5], nb = bi[6];
    const bool blk_local = LOCAL || (bi[8] != 0);
    const int64_t C = p.C;
    const int c0 = blockIdx.y * p.cs;
    const int nc = (int)min((int64_t)p.cs, C - c0);    // valid conformations of this chunk
    const bool tma = p.use_tma != 0;

    // ---- phase 0 ---------------------------------------------------------------------------------------------------
    if (tma) {
        if (tid == 0) e_mbar_init(mbar, 1);
        __syncthreads();
        if (warp == 0) {
            if (lane == 0) e_mbar_expect_tx(mbar, (uint32_t)(na * nc * 12));
            __syncwarp();
            for (int al = lane; al < na; al += 32)
                bulk_g2s(xs + al * kRow, p.xyz + ((int64_t)(a0 + al) * C + c0) * 3, (uint32_t)(nc * 12), mbar);
        }
    } else {
        const int per = nc * 3;
        for (int j = tid; j < na * per; j += NT) {
            const int al = j / per, rem = j - al * per;
            xs[al * kRow + rem] = __ldg(p.xyz + ((int64_t)(a0 + al) * C + c0) * 3 + rem);
        }
    }
    {
        const int4* src = reinterpret_cast<const int4*>(p.prog + prog0);     // blocks start at multiples of 4 words
        int4* dst = reinterpret_cast<int4*>(prog);
        for (int j = tid; j < (n_words >> 2); j += NT) dst[j] = __ldg(src + j);
    }
    int toff[4];
#pragma unroll
    for (int l = 0; l < 4; ++l) toff[l] = __ldg(p.mol_toff + (int64_t)l * (p.n_mol + 1) + mol);
    const float *pk2, *peq2, *pk3, *peq3, *pk4, *pk4i;
    if (LOCAL) {
        int cnt[4];
#pragma unroll
        for (int l = 0; l < 4; ++l) cnt[l] = __ldg(p.mol_toff + (int64_t)l * (p.n_mol + 1) + mol + 1) - toff[l];
        float* q = prm;
        pk2 = q;  for (int j = tid; j < cnt[0]; j += NT) q[j] = __ldg(p.k[0] + toff[0] + j);   q += cnt[0];
        peq2 = q; for (int j = tid; j < cnt[0]; j += NT) q[j] = __ldg(p.eq[0] + toff[0] + j);  q += cnt[0];
        pk3 = q;  for (int j = tid; j < cnt[1]; j += NT) q[j] = __ldg(p.k[1] + toff[1] + j);   q += cnt[1];
        peq3 = q; for (int j = tid; j < cnt[1]; j += NT) q[j] = __ldg(p.eq[1] + toff[1] + j);  q += cnt[1];
        pk4 = q;  for (int j = tid; j < cnt[2] * p.P4; j += NT) q[j] = __ldg(p.k[2] + (int64_t)toff[2] * p.P4 + j);   q += cnt[2] * p.P4;
        pk4i = q; for (int j = tid; j < cnt[3] * p.P4i; j += NT) q[j] = __ldg(p.k[3] + (int64_t)toff[3] * p.P4i + j);
    } else {
        pk2 = p.k[0] + toff[0]; peq2 = p.eq[0] + toff[0]; pk3 = p.k[1] + toff[1]; peq3 = p.eq[1] + toff[1];
        pk4 = p.k[2] + (int64_t)toff[2] * p.P4; pk4i = p.k[3] + (int64_t)toff[3] * p.P4i;
    }
    {   // every warp clears its own accumulator
        const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
        float4* a4 = reinterpret_cast<float4*>(acc + warp * S96);
        for (int j = lane; j < n_slots * (kRow / 4); j += 32) a4[j] = z;
    }
    __syncthreads();
    if (tma) e_mbar_wait(mbar, 0);

    // ---- phase 1: lane = conformation, warp = one part of the block's program -----------------------------------------
    float e0 = 0.f, e1 = 0.f, e2 = 0.f, e3 = 0.f;
    if (lane < nc) {
        const int* part = prog + prog[warp];
        const int n_prop = part[0], n_imp = part[1], n_ang = part[2], n_bond = part[3];
        const int* pw = part + 4;
        LaneCtx c;
        c.xl = reinterpret_cast<const char*>(xs) + lane * 12;
        c.accw = reinterpret_cast<char*>(acc + warp * S96) + lane * 12;
        c.C3 = C * 3; c.C = C; c.col = (int64_t)c0 + lane; c.na = na; c.blk_local = blk_local;
        c.halo = p.xyz + ((int64_t)a0 * C + c0 + lane) * 3;
        const int mask = p.term_mask;

        if (mask & 4) e2 = torsion_dispatch<LOCAL, MAT>(p.P4, n_prop, pw, c, pk4, toff[2], p.x[2], p.e[2]);
        pw += 5 * n_prop;
        if (mask & 8) e3 = torsion_dispatch<LOCAL, MAT>(p.P4i, n_imp, pw, c, pk4i, toff[3], p.x[3], p.e[3]);
        pw += 5 * n_imp;
        if (mask & 2) {
            for (int i = 0; i < n_ang; ++i, pw += 4) {
                const int hdr = pw[0];
                float *r0, *r1, *r2;
                const V3 x0 = rec_pos<LOCAL>(pw[1], c, r0);
                const V3 x1 = rec_pos<LOCAL>(pw[2], c, r1);
                const V3 x2 = rec_pos<LOCAL>(pw[3], c, r2);
                float th; V3 g0, g2;
                angle_geom(x0, x1, x2, th, g0, g2);
                const int tl = hdr >> 1;
                const float d = th - peq3[tl], f = pk3[tl] * d;
                acc3<LOCAL>(r0, f, g0);
                acc3<LOCAL>(r1, -f, g0 + g2);
                acc3<LOCAL>(r2, f, g2);
                if (LOCAL || (hdr & 1)) {
                    const float E = 0.5f * f * d;
                    e1 += E;
                    if (MAT) {
                        const int64_t row = (int64_t)(toff[1] + tl) * C + c.col;
                        if (p.x[1]) p.x[1][row] = th;
                        if (p.e[1]) p.e[1][row] = E;
                    }
                }
            }
        } else {
            pw += 4 * n_ang;
        }
        if (mask & 1) {
            for (int i = 0; i < n_bond; ++i, pw += 3) {
                const int hdr = pw[0];
                float *r0, *r1;
                const V3 x0 = rec_pos<LOCAL>(pw[1], c, r0);
                const V3 x1 = rec_pos<LOCAL>(pw[2], c, r1);
                float r_; V3 g0;
                bond_geom(x0, x1, r_, g0);
                const int tl = hdr >> 1;
                const float d = r_ - peq2[tl], f = pk2[tl] * d;
                acc3<LOCAL>(r0, f, g0);
                acc3<LOCAL>(r1, -f, g0);
                if (LOCAL || (hdr & 1)) {
                    const float E = 0.5f * f * d;
                    e0 += E;
                    if (MAT) {
                        const int64_t row = (int64_t)(toff[0] + tl) * C + c.col;
                        if (p.x[0]) p.x[0][row] = r_;
                        if (p.e[0]) p.e[0][row] = E;
                    }
                }
            }
        }
    }
    ered[(warp * 4 + 0) * 32 + lane] = e0;
    ered[(warp * 4 + 1) * 32 + lane] = e1;
    ered[(warp * 4 + 2) * 32 + lane] = e2;
    ered[(warp * 4 + 3) * 32 + lane] = e3;
    __syncthreads();

    // ---- phase 2: every atom's row = sum over the parts that touch it, in part order; the rows leave through the
    //      coordinate buffer (TMA store) or directly (coalesced stores) ------------------------------------------------
    const int* pull_ptr = prog + (W + 1);
    const int* pull = pull_ptr + (na + 1);             // part << 16 | float offset of the atom's slot in that part
    if (tma) {
        float4* o4 = reinterpret_cast<float4*>(xs);
        for (int j = tid; j < na * (kRow / 4); j += NT) {
            const int al = j / (kRow / 4), q = j - al * (kRow / 4);
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            for (int e = pull_ptr[al]; e < pull_ptr[al + 1]; ++e) {
                const int m = pull[e];
                const float4 u = reinterpret_cast<const float4*>(acc + (m >> 16) * S96 + (m & 0xffff))[q];
                v.x += u.x; v.y += u.y; v.z += u.z; v.w += u.w;
            }
            o4[j] = v;
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncthreads();
        if (warp == 0) {
            for (int al = lane; al < na; al += 32)
                bulk_s2g(p.gradient + ((int64_t)(a0 + al) * C + c0) * 3, xs + al * kRow, (uint32_t)(nc * 12));
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        }
    } else {
        const int per = nc * 3;
        for (int j = tid; j < na * per; j += NT) {
            const int al = j / per, rem = j - al * per;
            float v = 0.f;
            for (int e = pull_ptr[al]; e < pull_ptr[al + 1]; ++e) {
                const int m = pull[e];
                v += acc[(m >> 16) * S96 + (m & 0xffff) + rem];
            }
            p.gradient[((int64_t)(a0 + al) * C + c0) * 3 + rem] = v;
        }
    }

    // energies: lane c of warp 0 sums its conformation's partials in warp order
    if (tid < nc) {
        float s[4];
#pragma unroll
        for (int l = 0; l < 4; ++l) {
            float v = ered[l * 32 + tid];
            for (int w = 1; w < W; ++w) v += ered[(w * 4 + l) * 32 + tid];
            s[l] = v;
        }
        const int64_t conf = (int64_t)c0 + tid;
        if (nb == 1) {
#pragma unroll
            for (int l = 0; l < 4; ++l) p.term_energy[((int64_t)l * p.n_mol + mol) * C + conf] = s[l];
            p.energy[(int64_t)mol * C + conf] = ((s[0] + s[1]) + s[2]) + s[3];
        } else {
#pragma unroll
            for (int l = 0; l < 4; ++l) p.partial[((int64_t)b * 4 + l) * C + conf] = s[l];
        }
    }
    if (tma && warp == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");   // smem must outlive the stores
}

// molecules that span several blocks: sum the per-block partials in block order
__global__ void energy_finalize_kernel(EnergyParams p, int64_t n_blocks) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_blocks * p.C) return;
    const int64_t b = i / p.C, conf = i % p.C;
    const int32_t* bi = p.blk + b * kBlkStride;
    if (bi[6] <= 1 || bi[7] != 0) return;
    const int nb = bi[6], mol = bi[2];
    float tot = 0.f;
    for (int l = 0; l < 4; ++l) {
        float s = 0.f;
        for (int j = 0; j < nb; ++j) s += p.partial[((b + j) * 4 + l) * p.C + conf];
        p.term_energy[((int64_t)l * p.n_mol + mol) * p.C + conf] = s;
        tot += s;
    }
    p.energy[(int64_t)mol * p.C + conf] = tot;
}

// ------------------------------------------------------------------------------------------------- VJP --------

struct VjpParams {
    const float* xyz; const float* gE; const float* gG; const int32_t* tuple_mol;
    const int32_t* idx[4]; int64_t T[4]; int64_t Toff[5];
    const float* k[4]; const float* eq[2];
    int P4, P4i;
    float* dk[4]; float* deq[2];
    int64_t C;
};

// one warp per tuple; lanes stride over conformations; fixed-order shuffle reduction
__global__ void __launch_bounds__(256) energy_vjp_kernel(VjpParams p) {
    const int64_t w = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (w >= p.Toff[4]) return;
    int level = 0;
    while (level < 3 && w >= p.Toff[level + 1]) ++level;
    const int64_t t = w - p.Toff[level];
    const int n_seq = level == 0 ? 2 : (level == 1 ? 3 : 4);
    const int32_t* ix = p.idx[level] + t * n_seq;
    int a[4];
    for (int i = 0; i < n_seq; ++i) a[i] = __ldg(ix + i);
    const int mol = __ldg(p.tuple_mol + w);
    const int64_t C = p.C;

    float acc_k[kMaxPeriodicity] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    float acc_eq = 0.f;
    const int P = level == 2 ? p.P4 : (level == 3 ? p.P4i : 1);
    float kk = 0.f, eq = 0.f;
    if (level < 2) { kk = __ldg(p.k[level] + t); eq = __ldg(p.eq[level] + t); }

    for (int64_t conf = lane; conf < C; conf += 32) {
        const float* base = p.xyz + conf * 3;
        V3 x[4], gg[4];
        for (int i = 0; i < n_seq; ++i) x[i] = load3(base + (int64_t)a[i] * C * 3);
        const float ge = p.gE ? __ldg(p.gE + (int64_t)mol * C + conf) : 0.f;
        float cphi = 1.f, sphi = 0.f, d = 0.f;
        if (level == 0) {
            float r_; V3 g0; bond_geom(x[0], x[1], r_, g0);
            gg[0] = g0; gg[1] = (-1.f) * g0; d = r_ - eq;
        } else if (level == 1) {
            float th; V3 g0, g2; angle_geom(x[0], x[1], x[2], th, g0, g2);
            gg[0] = g0; gg[2] = g2; gg[1] = (-1.f) * (g0 + g2); d = th - eq;
        } else {
            dihedral_geom(x[0], x[1], x[2], x[3], cphi, sphi, gg);
        }
        float s = 0.f;   // directional derivative of the internal coordinate along dL/d(gradient)
        if (p.gG) {
            for (int i = 0; i < n_seq; ++i) {
                const V3 u = load3(p.gG + ((int64_t)a[i] * C + conf) * 3);
                s += dot(u, gg[i]);
            }
        }
        if (level < 2) {
            acc_k[0] += ge * 0.5f * d * d + d * s;
            acc_eq += -kk * (d * ge + s);
        } else {
            float cn = cphi, sn = sphi;
            for (int n = 1; n <= P; ++n) {
                acc_k[n - 1] += ge * cn - (float)n * sn * s;
                const float c2 = cn * cphi - sn * sphi;
                sn = sn * cphi + cn * sphi;
                cn = c2;
            }
        }
    }
    for (int n = 0; n < P; ++n) acc_k[n] = warp_sum(acc_k[n]);
    acc_eq = warp_sum(acc_eq);
    if (lane == 0) {
        if (level < 2) {
            p.dk[level][t] = acc_k[0];
            p.deq[level][t] = acc_eq;
        } else {
            for (int n = 0; n < P; ++n) p.dk[level][t * P + n] = acc_k[n];
        }
    }
}

template <bool LOCAL, bool MAT>
static int launch_energy2(const EnergyParams& p, int64_t n_blocks, int W, size_t smem, cudaStream_t st) {
    // the opt-in shared-memory size is a per-device attribute of the kernel: keep the largest value set per device
    static std::mutex mu;
    static size_t configured[64] = {0};
    int dev = 0;
    GRAPPA_CUDA(cudaGetDevice(&dev));
    if (smem > 48 * 1024) {
        std::lock_guard<std::mutex> lock(mu);
        if (dev < 0 || dev >= 64 || smem > configured[dev]) {
            GRAPPA_CUDA(cudaFuncSetAttribute(energy_fwd_kernel<LOCAL, MAT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            if (dev >= 0 && dev < 64) configured[dev] = smem;
        }
    }
    dim3 grid((unsigned)n_blocks, (unsigned)ceil_div(p.C, p.cs));
    energy_fwd_kernel<LOCAL, MAT><<<grid, 32 * W, smem, st>>>(p);
    GRAPPA_CHECK_LAUNCH();
    return 0;
}

}  // namespace grappa

extern "C" int grappa_energy_fwd(int64_t n_atoms, int64_t n_mol, int64_t C, const float* xyz,
                                 int64_t n_blocks, const int32_t* blk, const int32_t* prog, const int32_t* mol_tuple_off,
                                 const int64_t* plan_sizes, int term_mask,
                                 const float* k2, const float* eq2, const float* k3, const float* eq3,
                                 const float* k4, int P4, const float* k4i, int P4i,
                                 float* energy, float* term_energy, float* gradient,
                                 float* x2, float* x3, float* x4, float* x4i, float* e2, float* e3, float* e4, float* e4i,
                                 float* partial, void* stream) {
    GRAPPA_CHECK_ARG(xyz && blk && prog && mol_tuple_off && plan_sizes, "null plan/xyz pointer");
    GRAPPA_CHECK_ARG(energy && term_energy && gradient, "null output pointer");
    GRAPPA_CHECK_ARG(P4 >= 0 && P4 <= kMaxPeriodicity && P4i >= 0 && P4i <= kMaxPeriodicity, "periodicity must be <= 6");
    const int64_t max_words = plan_sizes[2], max_slots = plan_sizes[3], max_atoms = plan_sizes[4];
    const int any_multi = (int)plan_sizes[5], W = (int)plan_sizes[6], any_nonlocal = (int)plan_sizes[11];
    GRAPPA_CHECK_ARG(!any_multi || partial, "partial scratch required when a molecule spans several blocks");
    GRAPPA_CHECK_ARG(n_blocks < 2147483647LL && C < 65535LL * 32, "grid too large");
    GRAPPA_CHECK_ARG(n_blocks == plan_sizes[0], "n_blocks does not match the plan");
    if (n_atoms == 0 || C == 0 || n_blocks == 0) return 0;
    GRAPPA_CHECK_ARG(max_atoms > 0 && max_slots > 0 && max_words > 0, "empty plan sizes");
    EnergyParams p;
    p.xyz = xyz; p.blk = blk; p.prog = prog; p.mol_toff = mol_tuple_off;
    p.k[0] = k2; p.k[1] = k3; p.k[2] = k4; p.k[3] = k4i; p.eq[0] = eq2; p.eq[1] = eq3;
    p.P4 = P4; p.P4i = P4i; p.energy = energy; p.term_energy = term_energy; p.gradient = gradient;
    p.x[0] = x2; p.x[1] = x3; p.x[2] = x4; p.x[3] = x4i; p.e[0] = e2; p.e[1] = e3; p.e[2] = e4; p.e[3] = e4i;
    p.partial = partial; p.n_mol = n_mol; p.C = C; p.term_mask = term_mask;
    p.max_block_atoms = (int)max_atoms; p.max_slots = (int)max_slots; p.max_words = (int)((max_words + 3) / 4 * 4);
    const bool local = !any_nonlocal;
    const int64_t prm = local ? 2 * plan_sizes[7] + 2 * plan_sizes[8] + (int64_t)P4 * plan_sizes[9] + (int64_t)P4i * plan_sizes[10] : 0;
    p.prm_floats = (int)((prm + 3) / 4 * 4);
    // conformation chunks: as few as possible, balanced; rows move through the TMA engine when every row segment is
    // 16-byte aligned (C % 4 == 0, chunk size % 4 == 0, aligned base pointers), else through coalesced loads / stores
    const int64_t n_chunks = ceil_div(C, 32);
    int cs = (int)ceil_div(C, n_chunks);
    const bool aligned = (C % 4 == 0) && (reinterpret_cast<uintptr_t>(xyz) % 16 == 0) && (reinterpret_cast<uintptr_t>(gradient) % 16 == 0);
    if (aligned) cs = (cs + 3) / 4 * 4;
    static int tma_override = -1;
    if (tma_override < 0) { const char* e = getenv("GRAPPA_ENERGY_TMA"); tma_override = e ? (atoi(e) ? 1 : 0) : 2; }
    p.cs = cs;
    p.use_tma = aligned && tma_override != 0;
    const size_t smem = sizeof(float) * ((size_t)max_atoms * kRow + (size_t)W * max_slots * kRow + (size_t)W * 128 + p.prm_floats + p.max_words) + 16;
    GRAPPA_CHECK_ARG(smem <= 227 * 1024, "atom block too large for shared memory; rebuild the plan with fewer block_atoms or parts");
    cudaStream_t st = (cudaStream_t)stream;
    const bool mat = x2 || x3 || x4 || x4i || e2 || e3 || e4 || e4i;
    GRAPPA_CHECK_ARG(W >= 1 && W <= kMaxParts, "the plan's n_parts must be in [1, 16]");
    int rc;
    if (local) rc = mat ? launch_energy2<true, true>(p, n_blocks, W, smem, st) : launch_energy2<true, false>(p, n_blocks, W, smem, st);
    else rc = mat ? launch_energy2<false, true>(p, n_blocks, W, smem, st) : launch_energy2<false, false>(p, n_blocks, W, smem, st);
    if (rc) return rc;
    if (any_multi) {
        const int64_t n = n_blocks * C;
        energy_finalize_kernel<<<(unsigned)ceil_div(n, 256), 256, 0, st>>>(p, n_blocks);
        GRAPPA_CHECK_LAUNCH();
    }
    return 0;
}

extern "C" int grappa_energy_vjp(int64_t C, const float* xyz, const float* gE, const float* gG, const int32_t* tuple_mol,
                                 const int32_t* idx2, int64_t T2, const int32_t* idx3, int64_t T3,
                                 const int32_t* idx4, int64_t T4, const int32_t* idx4i, int64_t T4i,
                                 const float* k2, const float* eq2, const float* k3, const float* eq3,
                                 const float* k4, int P4, const float* k4i, int P4i,
                                 float* dk2, float* deq2, float* dk3, float* deq3, float* dk4, float* dk4i, void* stream) {
    GRAPPA_CHECK_ARG(xyz && tuple_mol, "null pointer");
    GRAPPA_CHECK_ARG(P4 >= 0 && P4 <= kMaxPeriodicity && P4i >= 0 && P4i <= kMaxPeriodicity, "periodicity must be <= 6");
    VjpParams p;
    p.xyz = xyz; p.gE = gE; p.gG = gG; p.tuple_mol = tuple_mol;
    p.idx[0] = idx2; p.idx[1] = idx3; p.idx[2] = idx4; p.idx[3] = idx4i;
    p.T[0] = T2; p.T[1] = T3; p.T[2] = T4; p.T[3] = T4i;
    p.Toff[0] = 0;
    for (int l = 0; l < 4; ++l) p.Toff[l + 1] = p.Toff[l] + p.T[l];
    p.k[0] = k2; p.k[1] = k3; p.k[2] = k4; p.k[3] = k4i; p.eq[0] = eq2; p.eq[1] = eq3;
    p.P4 = P4; p.P4i = P4i;
    p.dk[0] = dk2; p.dk[1] = dk3; p.dk[2] = dk4; p.dk[3] = dk4i; p.deq[0] = deq2; p.deq[1] = deq3;
    p.C = C;
    const int64_t n_warps = p.Toff[4];
    if (n_warps == 0) return 0;
    energy_vjp_kernel<<<(unsigned)ceil_div(n_warps * 32, 256), 256, 0, (cudaStream_t)stream>>>(p);
    GRAPPA_CHECK_LAUNCH();
    return 0;
}
