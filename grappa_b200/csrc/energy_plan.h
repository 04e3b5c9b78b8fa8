// Internal: the energy kernel's host plan as vectors (csrc/energy.cu builds it; grappa_energy_plan_build and
// grappa_plan_build copy it out).  Formats: csrc/energy.cu.
#pragma once
#include <stdint.h>
#include <vector>

namespace grappa {

struct EnergyPlanOut {
    std::vector<int32_t> blk, prog, toff, tuple_mol;
    int64_t sizes[12];
};

// fill = false: only `sizes` and `tuple_mol` are produced.
int build_energy_plan(int64_t n_atoms, int64_t n_mol, const int64_t* mol_atom_off,
                      const int64_t* idx2, int64_t T2, const int64_t* idx3, int64_t T3,
                      const int64_t* idx4, int64_t T4, const int64_t* idx4i, int64_t T4i,
                      int block_atoms, int n_parts, bool fill, EnergyPlanOut& out);

}  // namespace grappa
