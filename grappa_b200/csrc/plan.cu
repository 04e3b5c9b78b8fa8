// Host batch plan behind one entry point (grappa_plan_build / _free).  Everything index-shaped that
// the kernels need, derived from the graph's int64 idxs, its bond edges and the per-type node counts, laid out as
// sections of ONE packed int32 buffer (16-byte aligned sections), so that it reaches the device in one copy.
// The reference re-derives the same information on every call through advanced indexing and DGL's segment
// bookkeeping (models/interaction_parameters.py:184, models/internal_coordinates.py:31-94, models/energy.py:69).
//
// Sections, in buffer order (the order grappa_b200/plan.py exposes them under the same names):
//   per level n2, n3, n4, n4_improper:
//     idx        (T * n_seq)    tuple rows narrowed to int32
//     inc_ptr    (N + 1)        atom -> token-row incidence CSR (row = t * n_seq + position), stable in row order
//     inc_tok    (T * n_seq)
//     [levels with a positional-encoding pattern]
//     tab_idx    (T * n_seq)    (atom, pe) table row of every token = 2 * atom + pe(position)
//     tab_ptr    (2N + 1)       table row -> tokens CSR
//     tab_tok    (T * n_seq)
//   in_ptr (N + 1), in_nbr (E), out_ptr (N + 1), out_nbr (E)    bond edges by destination / by source
//   [with_energy]  e_blk, e_prog, e_mol_toff, tuple_mol             (csrc/energy.cu)
//   mol_atom_off (n_mol + 1), mol_off per level (n_mol + 1)
#include "common.cuh"
#include "energy_plan.h"
#include <thread>
#include <vector>

namespace grappa {

constexpr int kPlanSections = 4 * 6 + 4 + 4 + 1 + 4;     // 37 (absent sections have length 0)

struct PlanHandle {
    std::vector<int32_t> buf;
};

static inline int64_t pad4(int64_t n) { return (n + 3) / 4 * 4; }

// stable counting sort of n keys in [0, n_keys): ptr (n_keys + 1), order (n)
static void csr(int64_t n, int64_t n_keys, const int64_t* keys, int32_t* ptr, int32_t* order, std::vector<int32_t>& fill) {
    for (int64_t k = 0; k <= n_keys; ++k) ptr[k] = 0;
    for (int64_t i = 0; i < n; ++i) ptr[keys[i] + 1]++;
    for (int64_t k = 0; k < n_keys; ++k) ptr[k + 1] += ptr[k];
    fill.assign(ptr, ptr + n_keys);
    for (int64_t i = 0; i < n; ++i) order[fill[keys[i]]++] = (int32_t)i;
}

}  // namespace grappa

using namespace grappa;

extern "C" int grappa_plan_build(int64_t n_atoms, int64_t n_mol, const int64_t* mol_atom_off,
                                 const int64_t* const* idx /*[4]*/, const int64_t* T /*[4]*/,
                                 const int64_t* const* level_counts /*[4]: (n_mol) tuples per molecule, or NULL*/,
                                 const int32_t* const* pe_pattern /*[4]: (n_seq) or NULL*/,
                                 int64_t n_edges, const int64_t* edge_src, const int64_t* edge_dst,
                                 int with_energy, int block_atoms, int n_parts,
                                 int64_t* n_int32_out, int64_t* sections_out /*[2 * 37]*/, int64_t* energy_sizes_out /*[12]*/,
                                 void** handle_out) {
    GRAPPA_CHECK_ARG(n_atoms >= 0 && n_mol >= 0 && n_edges >= 0, "bad sizes");
    GRAPPA_CHECK_ARG(mol_atom_off && idx && T && level_counts && pe_pattern && n_int32_out && sections_out && handle_out, "null pointer");
    GRAPPA_CHECK_ARG(n_edges == 0 || (edge_src && edge_dst), "null edge arrays");
    GRAPPA_CHECK_ARG(2 * n_atoms + 1 < 2147483647LL && n_edges < 2147483647LL, "graph too large for int32 offsets");
    const int n_seq[4] = {2, 3, 4, 4};
    for (int l = 0; l < 4; ++l) {
        GRAPPA_CHECK_ARG(T[l] >= 0 && (T[l] == 0 || idx[l]) && T[l] * n_seq[l] < 2147483647LL, "bad tuple arrays");
        const int64_t n = T[l] * n_seq[l];
        int64_t mx = -1, mn = 0;
        for (int64_t i = 0; i < n; ++i) { mx = std::max(mx, idx[l][i]); mn = std::min(mn, idx[l][i]); }
        if (mx >= n_atoms || mn < 0) {
            // wording of models/grappa.py:179-185 (the Python wrapper turns this into an AssertionError)
            snprintf(g_err, sizeof(g_err), "Encountered idxs up to %lld at the level %d, but there are only %lld atom-level-nodes in the graph",
                     (long long)(mn < 0 ? mn : mx), l, (long long)n_atoms);
            return 4;
        }
    }
    for (int64_t e = 0; e < n_edges; ++e)
        GRAPPA_CHECK_ARG(edge_src[e] >= 0 && edge_src[e] < n_atoms && edge_dst[e] >= 0 && edge_dst[e] < n_atoms, "bond edge out of range");

    EnergyPlanOut eo;
    if (with_energy) {
        if (int rc = build_energy_plan(n_atoms, n_mol, mol_atom_off, idx[0], T[0], idx[1], T[1], idx[2], T[2], idx[3], T[3],
                                       block_atoms, n_parts, true, eo)) return rc;
        if (energy_sizes_out) for (int i = 0; i < 12; ++i) energy_sizes_out[i] = eo.sizes[i];
    }

    // ---- section table
    int64_t len[kPlanSections];
    int s = 0;
    for (int l = 0; l < 4; ++l) {
        const int64_t n = T[l] * n_seq[l];
        const bool pe = pe_pattern[l] != nullptr;
        len[s++] = n; len[s++] = n_atoms + 1; len[s++] = n;
        len[s++] = pe ? n : 0; len[s++] = pe ? 2 * n_atoms + 1 : 0; len[s++] = pe ? n : 0;
    }
    len[s++] = n_atoms + 1; len[s++] = n_edges; len[s++] = n_atoms + 1; len[s++] = n_edges;
    len[s++] = with_energy ? (int64_t)eo.blk.size() : 0;
    len[s++] = with_energy ? (int64_t)eo.prog.size() : 0;
    len[s++] = with_energy ? (int64_t)eo.toff.size() : 0;
    len[s++] = with_energy ? (int64_t)eo.tuple_mol.size() : 0;
    len[s++] = n_mol + 1;
    for (int l = 0; l < 4; ++l) len[s++] = n_mol + 1;
    int64_t total = 0;
    for (int i = 0; i < kPlanSections; ++i) {
        sections_out[2 * i] = total;
        sections_out[2 * i + 1] = len[i];
        total += pad4(len[i]);
    }
    total = std::max<int64_t>(total, 4);
    GRAPPA_CHECK_ARG(total < 2147483647LL, "plan too large");

    PlanHandle* h = new PlanHandle();
    h->buf.assign((size_t)total, 0);
    int32_t* B = h->buf.data();

    // the four levels and the edge tables are independent: one host thread each for large graphs
    auto level_job = [&](int l) {
        std::vector<int32_t> fill;
        std::vector<int64_t> keys;
        const int64_t n = T[l] * n_seq[l];
        int si = 6 * l;
        int32_t* o_idx = B + sections_out[2 * si]; ++si;
        int32_t* o_ptr = B + sections_out[2 * si]; ++si;
        int32_t* o_tok = B + sections_out[2 * si]; ++si;
        for (int64_t i = 0; i < n; ++i) o_idx[i] = (int32_t)idx[l][i];
        csr(n, n_atoms, idx[l], o_ptr, o_tok, fill);
        if (pe_pattern[l]) {
            int32_t* t_idx = B + sections_out[2 * si]; ++si;
            int32_t* t_ptr = B + sections_out[2 * si]; ++si;
            int32_t* t_tok = B + sections_out[2 * si]; ++si;
            keys.resize((size_t)n);
            for (int64_t i = 0; i < n; ++i) {
                keys[i] = 2 * idx[l][i] + pe_pattern[l][i % n_seq[l]];
                t_idx[i] = (int32_t)keys[i];
            }
            csr(n, 2 * n_atoms, keys.data(), t_ptr, t_tok, fill);
        }
    };
    auto edge_job = [&]() {
        std::vector<int32_t> fill;
        std::vector<int32_t> order((size_t)n_edges);
        int si = 24;
        int32_t* in_ptr = B + sections_out[2 * si]; ++si;
        int32_t* in_nbr = B + sections_out[2 * si]; ++si;
        csr(n_edges, n_atoms, edge_dst, in_ptr, order.data(), fill);
        for (int64_t e = 0; e < n_edges; ++e) in_nbr[e] = (int32_t)edge_src[order[e]];
        int32_t* out_ptr = B + sections_out[2 * si]; ++si;
        int32_t* out_nbr = B + sections_out[2 * si]; ++si;
        csr(n_edges, n_atoms, edge_src, out_ptr, order.data(), fill);
        for (int64_t e = 0; e < n_edges; ++e) out_nbr[e] = (int32_t)edge_dst[order[e]];
    };
    const bool threaded = (T[0] * 2 + T[1] * 3 + T[2] * 4 + T[3] * 4) > 20000;     // small batches: thread start-up costs more
    if (threaded) {
        std::thread th[4];
        for (int l = 1; l < 4; ++l) th[l] = std::thread(level_job, l);
        th[0] = std::thread(edge_job);
        level_job(0);
        for (int l = 0; l < 4; ++l) th[l].join();
    } else {
        for (int l = 0; l < 4; ++l) level_job(l);
        edge_job();
    }
    s = 28;
    if (with_energy) {
        std::copy(eo.blk.begin(), eo.blk.end(), B + sections_out[2 * s]); ++s;
        std::copy(eo.prog.begin(), eo.prog.end(), B + sections_out[2 * s]); ++s;
        std::copy(eo.toff.begin(), eo.toff.end(), B + sections_out[2 * s]); ++s;
        std::copy(eo.tuple_mol.begin(), eo.tuple_mol.end(), B + sections_out[2 * s]); ++s;
    } else {
        s += 4;
    }
    {
        int32_t* o = B + sections_out[2 * s]; ++s;
        for (int64_t m = 0; m <= n_mol; ++m) o[m] = (int32_t)mol_atom_off[m];
    }
    for (int l = 0; l < 4; ++l) {
        int32_t* o = B + sections_out[2 * s]; ++s;
        o[0] = 0;
        for (int64_t m = 0; m < n_mol; ++m) o[m + 1] = o[m] + (level_counts[l] ? (int32_t)level_counts[l][m] : 0);
    }
    *n_int32_out = total;
    *handle_out = h;
    return 0;
}

extern "C" int grappa_plan_fill(void* handle, int32_t* out, int64_t n_int32) {
    GRAPPA_CHECK_ARG(handle && out, "null pointer");
    PlanHandle* h = static_cast<PlanHandle*>(handle);
    GRAPPA_CHECK_ARG(n_int32 == (int64_t)h->buf.size(), "buffer size does not match the plan");
    std::copy(h->buf.begin(), h->buf.end(), out);
    return 0;
}

extern "C" void grappa_plan_free(void* handle) { delete static_cast<PlanHandle*>(handle); }
