// K4: fused MM energy + analytic dE/dxyz, K5: its VJP w.r.t. the MM parameters.  See include/grappa_b200.h.
//
// Data layout in HBM: xyz / gradient are (N, C, 3) fp32, atom-major, conformation-minor (the reference's layout,
// internal_coordinates.py:31).  One CTA evaluates (atom block, chunk of CC conformations):
//   phase 1  one thread per (tuple record, conformation): gathers the 2-4 atom positions (L1/L2-resident: a block's
//            coordinates are re-read ~19x), computes the internal coordinate, the energy and the per-atom gradient
//            vectors, and writes the gradient vectors of the atoms that belong to this block into a shared-memory
//            staging array (one slot per (record, role)).
//   phase 2  one thread per (atom, conformation): sums the atom's slots in the fixed order of the host plan and
//            writes gradient[a, c, :] once.  No atomics anywhere -> bit-reproducible forces.
//   energies per-thread partial sums -> fixed-order shared-memory reduction -> (molecule, conformation) row.
// Algorithmic HBM bytes per (molecule, conformation) with A atoms: 12A (xyz) + 12A (gradient) + 20 (energies).
#include "common.cuh"
#include "geometry.cuh"
#include <vector>
#include <algorithm>

namespace grappa {

// ------------------------------------------------------------------------------------------------ host plan ----

constexpr int kBlkStride = 12;   // ints per block row: a0, na, mol, rec0, n_rec, n_slots, n_blocks_of_mol, j, n_rec per level[4]

struct LevelIdx {
    const int64_t* idx;
    int64_t T;
    int n_seq;
};

}  // namespace grappa

using namespace grappa;

extern "C" int grappa_energy_plan_build(int64_t n_atoms, int64_t n_mol, const int64_t* mol_atom_off,
                                        const int64_t* idx2, int64_t T2, const int64_t* idx3, int64_t T3,
                                        const int64_t* idx4, int64_t T4, const int64_t* idx4i, int64_t T4i,
                                        int block_atoms, int64_t* sizes_out,
                                        int32_t* blk, int32_t* rec_atoms, int32_t* rec_meta, int32_t* rec_slot,
                                        int32_t* pull_ptr, int32_t* pull_slot, int32_t* tuple_mol) {
    GRAPPA_CHECK_ARG(n_atoms >= 0 && n_mol >= 0 && block_atoms > 0, "bad sizes");
    GRAPPA_CHECK_ARG(mol_atom_off != nullptr && sizes_out != nullptr, "null pointer");
    GRAPPA_CHECK_ARG(mol_atom_off[0] == 0 && mol_atom_off[n_mol] == n_atoms, "mol_atom_off must span [0, n_atoms]");
    LevelIdx lv[4] = {{idx2, T2, 2}, {idx3, T3, 3}, {idx4, T4, 4}, {idx4i, T4i, 4}};

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

    // pass 1: count records per block, slots per block, incidences per atom
    std::vector<int64_t> rec_cnt(nB + 1, 0), slot_cnt(nB, 0), inc_cnt(n_atoms + 1, 0);
    int64_t tuple_base = 0;
    for (int l = 0; l < 4; ++l) {
        for (int64_t t = 0; t < lv[l].T; ++t) {
            const int64_t* a = lv[l].idx + t * lv[l].n_seq;
            int32_t tb[4]; int nt = 0;
            for (int i = 0; i < lv[l].n_seq; ++i) {
                if (a[i] < 0 || a[i] >= n_atoms) return fail("grappa_energy_plan_build: %s atom index %lld out of range [0,%lld)", "", (long long)a[i], (long long)n_atoms);
                int32_t b = atom_blk[a[i]];
                bool seen = false;
                for (int q = 0; q < nt; ++q) seen |= (tb[q] == b);
                if (!seen) tb[nt++] = b;
                slot_cnt[b]++; inc_cnt[a[i] + 1]++;
            }
            for (int q = 0; q < nt; ++q) rec_cnt[tb[q] + 1]++;
            if (tuple_mol) tuple_mol[tuple_base + t] = atom_mol[a[0]];
        }
        tuple_base += lv[l].T;
    }
    for (int64_t b = 0; b < nB; ++b) rec_cnt[b + 1] += rec_cnt[b];
    for (int64_t a = 0; a < n_atoms; ++a) inc_cnt[a + 1] += inc_cnt[a];
    const int64_t nR = rec_cnt[nB], nP = inc_cnt[n_atoms];
    int64_t max_slots = 0, max_rec = 0, max_atoms = 0, any_multi = 0;
    for (int64_t b = 0; b < nB; ++b) {
        max_slots = std::max(max_slots, slot_cnt[b]);
        max_rec = std::max(max_rec, rec_cnt[b + 1] - rec_cnt[b]);
        max_atoms = std::max<int64_t>(max_atoms, blocks[b].na);
        any_multi |= (blocks[b].nb > 1);
    }
    sizes_out[0] = nB; sizes_out[1] = nR; sizes_out[2] = nP; sizes_out[3] = max_slots;
    sizes_out[4] = max_rec; sizes_out[5] = max_atoms; sizes_out[6] = any_multi; sizes_out[7] = 0;
    if (blk == nullptr) return 0;   // size query only
    GRAPPA_CHECK_ARG(rec_atoms && rec_meta && rec_slot && pull_ptr && pull_slot, "null output array");
    GRAPPA_CHECK_ARG(max_slots < (1 << 30), "block too large");

    for (int64_t b = 0; b < nB; ++b) {
        int32_t* o = blk + b * kBlkStride;
        o[0] = blocks[b].a0; o[1] = blocks[b].na; o[2] = blocks[b].mol; o[3] = (int32_t)rec_cnt[b];
        o[4] = (int32_t)(rec_cnt[b + 1] - rec_cnt[b]); o[5] = (int32_t)slot_cnt[b];
        o[6] = blocks[b].nb; o[7] = blocks[b].j;
        o[8] = o[9] = o[10] = o[11] = 0;      // records per level (filled in pass 2)
    }
    for (int64_t a = 0; a <= n_atoms; ++a) pull_ptr[a] = (int32_t)inc_cnt[a];

    // pass 2: fill records (ordered by level, then tuple, within each block), slots, pull lists
    std::vector<int64_t> rec_fill(rec_cnt.begin(), rec_cnt.end() - 1), slot_fill(nB, 0), inc_fill(inc_cnt.begin(), inc_cnt.end() - 1);
    for (int l = 0; l < 4; ++l) {
        for (int64_t t = 0; t < lv[l].T; ++t) {
            const int64_t* a = lv[l].idx + t * lv[l].n_seq;
            int32_t tb[4]; int nt = 0;
            for (int i = 0; i < lv[l].n_seq; ++i) {
                int32_t b = atom_blk[a[i]];
                bool seen = false;
                for (int q = 0; q < nt; ++q) seen |= (tb[q] == b);
                if (!seen) tb[nt++] = b;
            }
            const int32_t owner = atom_blk[a[0]];
            for (int q = 0; q < nt; ++q) {
                const int32_t b = tb[q];
                const int64_t r = rec_fill[b]++;
                blk[(int64_t)b * kBlkStride + 8 + l]++;
                rec_meta[r * 2 + 0] = l | ((b == owner) ? 4 : 0);
                rec_meta[r * 2 + 1] = (int32_t)t;
                for (int i = 0; i < 4; ++i) {
                    if (i < lv[l].n_seq) {
                        rec_atoms[r * 4 + i] = (int32_t)a[i];
                        if (atom_blk[a[i]] == b) {
                            const int32_t s = (int32_t)slot_fill[b]++;
                            rec_slot[r * 4 + i] = s;
                            pull_slot[inc_fill[a[i]]++] = s;
                        } else {
                            rec_slot[r * 4 + i] = -1;
                        }
                    } else {
                        rec_atoms[r * 4 + i] = 0;
                        rec_slot[r * 4 + i] = -1;
                    }
                }
            }
        }
    }
    return 0;
}

// --------------------------------------------------------------------------------------------- device code ----

namespace grappa {

struct EnergyParams {
    const float* xyz;
    const int32_t* blk;
    const int32_t* rec_atoms;
    const int32_t* rec_meta;
    const int32_t* rec_slot;
    const int32_t* pull_ptr;
    const int32_t* pull_slot;
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
    int max_block_atoms;
};

__device__ __forceinline__ V3 load3(const float* __restrict__ p) { return v3(__ldg(p), __ldg(p + 1), __ldg(p + 2)); }

constexpr int kEnergyThreads = 256;
constexpr int kMaxPeriodicity = 6;

// torsion energy sum_n k_n cos(n phi) and dE/dphi = -sum_n n k_n sin(n phi) from (cos phi, sin phi)
__device__ __forceinline__ void torsion_terms(const float* __restrict__ k, int P, float c, float s, float& E, float& dEdphi) {
    float cn = c, sn = s;
    E = 0.f; dEdphi = 0.f;
    for (int n = 1; n <= P; ++n) {
        const float kn = __ldg(k + n - 1);
        E = fmaf(kn, cn, E);
        dEdphi = fmaf(-(float)n * kn, sn, dEdphi);
        const float c2 = cn * c - sn * s;
        sn = sn * c + cn * s;
        cn = c2;
    }
}

// Position of one atom for conformation c (chunk-local).  LOCAL: every atom of every record belongs to this CTA's
// block (true whenever no molecule spans several blocks), so positions always come from the shared-memory copy.
template <int CC, bool LOCAL>
__device__ __forceinline__ V3 atom_pos(const float* __restrict__ xs, const float* __restrict__ xyz, int a, int a0, int na,
                                       int c, int64_t C, int64_t c0) {
    const unsigned al = (unsigned)(a - a0);
    if (LOCAL || al < (unsigned)na) {
        const float* q = xs + (al * CC + c) * 3;
        return v3(q[0], q[1], q[2]);
    }
    const float* q = xyz + ((int64_t)a * C + c0 + c) * 3;
    return v3(__ldg(q), __ldg(q + 1), __ldg(q + 2));
}

__device__ __forceinline__ void put3(float* __restrict__ stage, int slot, int cc, int c, float x, float y, float z) {
    if (slot >= 0) { float* o = stage + (slot * cc + c) * 3; o[0] = x; o[1] = y; o[2] = z; }
}

template <int CC, bool LOCAL>
__global__ void __launch_bounds__(kEnergyThreads) energy_fwd_kernel(EnergyParams p) {
    extern __shared__ float smem[];
    float* red = smem;                                        // [4][kEnergyThreads]
    float* xs = smem + 4 * kEnergyThreads;                    // [block atom][CC][3]
    float* stage = xs + (size_t)p.max_block_atoms * CC * 3;   // [slot][CC][3]

    const int b = blockIdx.x;
    const int64_t c0 = (int64_t)blockIdx.y * CC;
    const int32_t* bi = p.blk + (int64_t)b * kBlkStride;
    const int a0 = bi[0], na = bi[1], mol = bi[2], r0 = bi[3], nb = bi[6];
    const int n_bond = bi[8], n_angle = bi[9], n_proper = bi[10], n_improper = bi[11];
    const int tid = threadIdx.x;
    const int64_t C = p.C;
    const int nc = (int)min((int64_t)CC, C - c0);             // valid conformations of this chunk

    // phase 0: the block's coordinates for this chunk, one coalesced pass (CC*3 contiguous floats per atom)
    for (int j = tid; j < na * CC * 3; j += kEnergyThreads) {
        const int al = j / (CC * 3), rem = j % (CC * 3);
        xs[j] = (rem < nc * 3) ? __ldg(p.xyz + ((int64_t)(a0 + al) * C + c0) * 3 + rem) : 0.f;
    }
    __syncthreads();

    // phase 1: one thread per (tuple record, conformation); records are sorted by level inside a block, so each level
    // gets its own loop (no per-item level dispatch).  Masked levels write zeros into their slots.
    float e0 = 0.f, e1 = 0.f, e2 = 0.f, e3 = 0.f;
    const bool on0 = p.term_mask & 1, on1 = p.term_mask & 2, on2 = p.term_mask & 4, on3 = p.term_mask & 8;

    // One flattened item list ordered by cost (torsions, then angles, then bonds): every pass of 256 items is of
    // (nearly) one kind, and the per-thread work is balanced to within one cheap item before the barrier.
    const int n_tor = n_proper + n_improper;
    for (int item = tid; item < (n_tor + n_angle + n_bond) * CC; item += kEnergyThreads) {
      const int rl_all = item / CC, c = item % CC;
      if (c >= nc) continue;
      if (rl_all >= n_tor + n_angle) {                                             // ---- bonds
        const int r = r0 + (rl_all - n_tor - n_angle);
        const int2 meta = __ldg(reinterpret_cast<const int2*>(p.rec_meta) + r);
        const int4 at = __ldg(reinterpret_cast<const int4*>(p.rec_atoms) + r);
        const int4 sl = __ldg(reinterpret_cast<const int4*>(p.rec_slot) + r);
        float f = 0.f, E = 0.f, r_ = 0.f; V3 g0 = v3(0.f, 0.f, 0.f);
        if (on0) {
            bond_geom(atom_pos<CC, LOCAL>(xs, p.xyz, at.x, a0, na, c, C, c0), atom_pos<CC, LOCAL>(xs, p.xyz, at.y, a0, na, c, C, c0), r_, g0);
            const float k = __ldg(p.k[0] + meta.y), d = r_ - __ldg(p.eq[0] + meta.y);
            f = k * d; E = 0.5f * f * d;
        }
        put3(stage, sl.x, CC, c, f * g0.x, f * g0.y, f * g0.z);
        put3(stage, sl.y, CC, c, -f * g0.x, -f * g0.y, -f * g0.z);
        if (on0 && (meta.x & 4)) {
            e0 += E;
            if (p.x[0]) p.x[0][(int64_t)meta.y * C + c0 + c] = r_;
            if (p.e[0]) p.e[0][(int64_t)meta.y * C + c0 + c] = E;
        }
      } else if (rl_all >= n_tor) {                                                // ---- angles
        const int r = r0 + n_bond + (rl_all - n_tor);
        const int2 meta = __ldg(reinterpret_cast<const int2*>(p.rec_meta) + r);
        const int4 at = __ldg(reinterpret_cast<const int4*>(p.rec_atoms) + r);
        const int4 sl = __ldg(reinterpret_cast<const int4*>(p.rec_slot) + r);
        float f = 0.f, E = 0.f, th = 0.f; V3 g0 = v3(0.f, 0.f, 0.f), g2 = g0;
        if (on1) {
            angle_geom(atom_pos<CC, LOCAL>(xs, p.xyz, at.x, a0, na, c, C, c0), atom_pos<CC, LOCAL>(xs, p.xyz, at.y, a0, na, c, C, c0),
                       atom_pos<CC, LOCAL>(xs, p.xyz, at.z, a0, na, c, C, c0), th, g0, g2);
            const float k = __ldg(p.k[1] + meta.y), d = th - __ldg(p.eq[1] + meta.y);
            f = k * d; E = 0.5f * f * d;
        }
        put3(stage, sl.x, CC, c, f * g0.x, f * g0.y, f * g0.z);
        put3(stage, sl.y, CC, c, -f * (g0.x + g2.x), -f * (g0.y + g2.y), -f * (g0.z + g2.z));
        put3(stage, sl.z, CC, c, f * g2.x, f * g2.y, f * g2.z);
        if (on1 && (meta.x & 4)) {
            e1 += E;
            if (p.x[1]) p.x[1][(int64_t)meta.y * C + c0 + c] = th;
            if (p.e[1]) p.e[1][(int64_t)meta.y * C + c0 + c] = E;
        }
      } else {                                                                     // ---- proper + improper torsions
        const int rl = rl_all;
        const int r = r0 + n_bond + n_angle + rl;
        const bool improper = rl >= n_proper;
        const int2 meta = __ldg(reinterpret_cast<const int2*>(p.rec_meta) + r);
        const int4 at = __ldg(reinterpret_cast<const int4*>(p.rec_atoms) + r);
        const int4 sl = __ldg(reinterpret_cast<const int4*>(p.rec_slot) + r);
        float dE = 0.f, E = 0.f, cphi = 1.f, sphi = 0.f;
        V3 gg[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) gg[i] = v3(0.f, 0.f, 0.f);
        const bool on = improper ? on3 : on2;
        if (on) {
            dihedral_geom(atom_pos<CC, LOCAL>(xs, p.xyz, at.x, a0, na, c, C, c0), atom_pos<CC, LOCAL>(xs, p.xyz, at.y, a0, na, c, C, c0),
                          atom_pos<CC, LOCAL>(xs, p.xyz, at.z, a0, na, c, C, c0), atom_pos<CC, LOCAL>(xs, p.xyz, at.w, a0, na, c, C, c0),
                          cphi, sphi, gg);
            const int P = improper ? p.P4i : p.P4;
            torsion_terms((improper ? p.k[3] : p.k[2]) + (int64_t)meta.y * P, P, cphi, sphi, E, dE);
        }
        put3(stage, sl.x, CC, c, dE * gg[0].x, dE * gg[0].y, dE * gg[0].z);
        put3(stage, sl.y, CC, c, dE * gg[1].x, dE * gg[1].y, dE * gg[1].z);
        put3(stage, sl.z, CC, c, dE * gg[2].x, dE * gg[2].y, dE * gg[2].z);
        put3(stage, sl.w, CC, c, dE * gg[3].x, dE * gg[3].y, dE * gg[3].z);
        if (on && (meta.x & 4)) {
            if (improper) e3 += E; else e2 += E;
            float* xo = improper ? p.x[3] : p.x[2];
            float* eo = improper ? p.e[3] : p.e[2];
            if (xo) xo[(int64_t)meta.y * C + c0 + c] = atan2f(sphi, cphi);
            if (eo) eo[(int64_t)meta.y * C + c0 + c] = E;
        }
      }
    }
    red[0 * kEnergyThreads + tid] = e0;
    red[1 * kEnergyThreads + tid] = e1;
    red[2 * kEnergyThreads + tid] = e2;
    red[3 * kEnergyThreads + tid] = e3;
    __syncthreads();

    // phase 2: pull-reduce the staged gradient vectors per (atom, conformation), fixed order
    for (int j = tid; j < na * CC; j += kEnergyThreads) {
        const int al = j / CC, c = j % CC;
        if (c >= nc) continue;
        const int a = a0 + al;
        const int q0 = __ldg(p.pull_ptr + a), q1 = __ldg(p.pull_ptr + a + 1);
        float gx = 0.f, gy = 0.f, gz = 0.f;
        for (int q = q0; q < q1; ++q) {
            const float* sp = stage + (__ldg(p.pull_slot + q) * CC + c) * 3;
            gx += sp[0]; gy += sp[1]; gz += sp[2];
        }
        float* o = p.gradient + ((int64_t)a * C + c0 + c) * 3;
        o[0] = gx; o[1] = gy; o[2] = gz;
    }

    // energies: thread j < 4*CC sums its (term, conformation) column over the CTA in thread order
    // (item % CC == tid % CC for every item of a thread because kEnergyThreads % CC == 0)
    float tsum = 0.f;
    if (tid < 4 * CC) {
        const int l = tid / CC, c = tid % CC;
        for (int i = c; i < kEnergyThreads; i += CC) tsum += red[l * kEnergyThreads + i];
    }
    __syncthreads();
    if (tid < 4 * CC) red[tid] = tsum;
    __syncthreads();
    if (tid < 4 * CC) {
        const int l = tid / CC, c = tid % CC;
        const int64_t conf = c0 + c;
        if (conf < C) {
            if (nb == 1) {
                p.term_energy[((int64_t)l * p.n_mol + mol) * C + conf] = tsum;
                if (l == 0) p.energy[(int64_t)mol * C + conf] = ((red[c] + red[CC + c]) + red[2 * CC + c]) + red[3 * CC + c];
            } else {
                p.partial[((int64_t)b * 4 + l) * C + conf] = tsum;
            }
        }
    }
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

template <int CC, bool LOCAL>
static int launch_energy2(const EnergyParams& p, int64_t n_blocks, int max_slots, cudaStream_t st) {
    const size_t smem = (size_t)(4 * kEnergyThreads + ((size_t)max_slots + p.max_block_atoms) * CC * 3) * sizeof(float);
    static size_t configured = 0;
    if (smem > 48 * 1024 && smem > configured) {
        GRAPPA_CUDA(cudaFuncSetAttribute(energy_fwd_kernel<CC, LOCAL>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured = smem;
    }
    dim3 grid((unsigned)n_blocks, (unsigned)ceil_div(p.C, CC));
    energy_fwd_kernel<CC, LOCAL><<<grid, kEnergyThreads, smem, st>>>(p);
    GRAPPA_CHECK_LAUNCH();
    return 0;
}

template <int CC>
static int launch_energy(const EnergyParams& p, int64_t n_blocks, int max_slots, bool local, cudaStream_t st) {
    return local ? launch_energy2<CC, true>(p, n_blocks, max_slots, st) : launch_energy2<CC, false>(p, n_blocks, max_slots, st);
}

}  // namespace grappa

static int g_energy_cc_override = -1;

extern "C" int grappa_energy_fwd(int64_t n_atoms, int64_t n_mol, int64_t C, const float* xyz,
                                 int64_t n_blocks, const int32_t* blk, const int32_t* rec_atoms, const int32_t* rec_meta,
                                 const int32_t* rec_slot, const int32_t* pull_ptr, const int32_t* pull_slot,
                                 int max_slots, int max_block_atoms, int any_multi_block, int term_mask,
                                 const float* k2, const float* eq2, const float* k3, const float* eq3,
                                 const float* k4, int P4, const float* k4i, int P4i,
                                 float* energy, float* term_energy, float* gradient,
                                 float* x2, float* x3, float* x4, float* x4i, float* e2, float* e3, float* e4, float* e4i,
                                 float* partial, void* stream) {
    GRAPPA_CHECK_ARG(xyz && blk && rec_atoms && rec_meta && rec_slot && pull_ptr && pull_slot, "null plan/xyz pointer");
    GRAPPA_CHECK_ARG(energy && term_energy && gradient, "null output pointer");
    GRAPPA_CHECK_ARG(P4 >= 0 && P4 <= kMaxPeriodicity && P4i >= 0 && P4i <= kMaxPeriodicity, "periodicity must be <= 6");
    GRAPPA_CHECK_ARG(!any_multi_block || partial, "partial scratch required when a molecule spans several blocks");
    GRAPPA_CHECK_ARG(n_blocks < 2147483647LL && ceil_div(C, 1) < 65535LL * 32, "grid too large");
    if (n_atoms == 0 || C == 0 || n_blocks == 0) return 0;
    EnergyParams p;
    p.xyz = xyz; p.blk = blk; p.rec_atoms = rec_atoms; p.rec_meta = rec_meta; p.rec_slot = rec_slot;
    p.pull_ptr = pull_ptr; p.pull_slot = pull_slot;
    p.k[0] = k2; p.k[1] = k3; p.k[2] = k4; p.k[3] = k4i; p.eq[0] = eq2; p.eq[1] = eq3;
    p.P4 = P4; p.P4i = P4i; p.energy = energy; p.term_energy = term_energy; p.gradient = gradient;
    p.x[0] = x2; p.x[1] = x3; p.x[2] = x4; p.x[3] = x4i; p.e[0] = e2; p.e[1] = e3; p.e[2] = e4; p.e[3] = e4i;
    p.partial = partial; p.n_mol = n_mol; p.C = C; p.term_mask = term_mask; p.max_block_atoms = max_block_atoms;
    cudaStream_t st = (cudaStream_t)stream;

    // conformations per CTA: the staging array holds max_slots * CC float3; keep it near 40 KB so that ~5 CTAs
    // (1280 threads) stay resident per SM, and never wider than the conformation count.
    int cc = 32;
    if (g_energy_cc_override < 0) {
        const char* e = getenv("GRAPPA_ENERGY_CC");
        g_energy_cc_override = e ? atoi(e) : 0;
    }
    if (g_energy_cc_override > 0) cc = g_energy_cc_override;
    else while (cc > 1 && ((int64_t)(max_slots + max_block_atoms) * cc * 12 > 40 * 1024 || cc / 2 >= C)) cc >>= 1;
    GRAPPA_CHECK_ARG(max_block_atoms > 0, "max_block_atoms must be positive");
    GRAPPA_CHECK_ARG((int64_t)(max_slots + max_block_atoms) * cc * 12 + 4096 <= 200 * 1024, "atom block too large for shared memory; rebuild the plan with fewer block_atoms");
    GRAPPA_CHECK_ARG(ceil_div(C, cc) <= 65535, "too many conformation chunks");
    int rc;
    switch (cc) {
        case 32: rc = launch_energy<32>(p, n_blocks, max_slots, !any_multi_block, st); break;
        case 16: rc = launch_energy<16>(p, n_blocks, max_slots, !any_multi_block, st); break;
        case 8: rc = launch_energy<8>(p, n_blocks, max_slots, !any_multi_block, st); break;
        case 4: rc = launch_energy<4>(p, n_blocks, max_slots, !any_multi_block, st); break;
        case 2: rc = launch_energy<2>(p, n_blocks, max_slots, !any_multi_block, st); break;
        case 1: rc = launch_energy<1>(p, n_blocks, max_slots, !any_multi_block, st); break;
        default: GRAPPA_CHECK_ARG(false, "GRAPPA_ENERGY_CC must be a power of two <= 32"); rc = 2;
    }
    if (rc) return rc;
    if (any_multi_block) {
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
