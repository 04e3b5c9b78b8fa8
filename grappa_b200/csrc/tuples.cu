// Host-side interaction-tuple enumeration: bonds -> angles and proper torsions, rows in the
// reference's order.  Replaces the Python loop nest of utils/tuple_indices.py:8-70 (get_idx_tuples) and the neighbour
// dictionary of :73-110 (get_neighbor_dict):
//   outer loop over atoms in the order in which they first appear in the flattened bond list (dict insertion order),
//   neighbour lists sorted ascending, angle (a1,a2,a3) kept when a1 < a3, proper (a4,a3,a2,a1) for a4 < a1, a4 != a2.
// The row ORDER is part of the contract (the model's outputs are rows aligned with these tuples).
#include "common.cuh"
#include <algorithm>
#include <vector>

using namespace grappa;

extern "C" int grappa_enumerate_tuples(int64_t n_bonds, const int64_t* bonds, int64_t n_atoms, int64_t* counts,
                                       int64_t* angles, int64_t* propers) {
    GRAPPA_CHECK_ARG(n_bonds >= 0 && n_atoms >= 0 && counts && (n_bonds == 0 || bonds), "bad arguments");
    GRAPPA_CHECK_ARG((angles == nullptr) == (propers == nullptr), "angles and propers must both be given or both be NULL");
    const bool fill = angles != nullptr;
    const int64_t cap_angles = counts[0], cap_propers = counts[1];

    std::vector<int64_t> ptr(n_atoms + 1, 0);
    for (int64_t b = 0; b < n_bonds; ++b) {
        const int64_t i = bonds[2 * b], j = bonds[2 * b + 1];
        if (i < 0 || i >= n_atoms || j < 0 || j >= n_atoms)
            return fail("grappa_enumerate_tuples:%s bond %lld has an atom outside [0,%lld)", "", (long long)b, (long long)n_atoms);
        if (i == j) return fail("grappa_enumerate_tuples: %s self-bond at row %lld", "Encountered", (long long)b);
        ptr[i + 1]++;
        ptr[j + 1]++;
    }
    for (int64_t a = 0; a < n_atoms; ++a) ptr[a + 1] += ptr[a];
    std::vector<int64_t> nbr(2 * n_bonds), cur(ptr.begin(), ptr.end() - 1), order;
    std::vector<char> seen(n_atoms, 0);
    order.reserve(n_atoms);
    for (int64_t b = 0; b < n_bonds; ++b) {
        for (int s = 0; s < 2; ++s) {
            const int64_t a = bonds[2 * b + s], o = bonds[2 * b + 1 - s];
            nbr[cur[a]++] = o;
            if (!seen[a]) { seen[a] = 1; order.push_back(a); }
        }
    }
    for (int64_t a = 0; a < n_atoms; ++a) std::sort(nbr.begin() + ptr[a], nbr.begin() + ptr[a + 1]);

    int64_t na = 0, np = 0;
    for (int64_t a1 : order) {
        for (int64_t e2 = ptr[a1]; e2 < ptr[a1 + 1]; ++e2) {
            const int64_t a2 = nbr[e2];
            for (int64_t e3 = ptr[a2]; e3 < ptr[a2 + 1]; ++e3) {
                const int64_t a3 = nbr[e3];
                if (a3 == a1) continue;
                if (a1 < a3) {
                    if (fill) {
                        if (na >= cap_angles) return fail("grappa_enumerate_tuples:%s angle capacity %lld too small", "", (long long)cap_angles);
                        int64_t* r = angles + 3 * na;
                        r[0] = a1; r[1] = a2; r[2] = a3;
                    }
                    ++na;
                }
                for (int64_t e4 = ptr[a3]; e4 < ptr[a3 + 1]; ++e4) {
                    const int64_t a4 = nbr[e4];
                    if (a4 >= a1) break;
                    if (a4 == a2) continue;
                    if (fill) {
                        if (np >= cap_propers) return fail("grappa_enumerate_tuples:%s proper capacity %lld too small", "", (long long)cap_propers);
                        int64_t* r = propers + 4 * np;
                        r[0] = a4; r[1] = a3; r[2] = a2; r[3] = a1;
                    }
                    ++np;
                }
            }
        }
    }
    counts[0] = na;
    counts[1] = np;
    return 0;
}
