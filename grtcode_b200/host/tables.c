/* Species tables (data of gas-optics/src/molecules.c:36-306, cfcs.c:40-103,
 * collision_induced_absorption.c:46-56). */
#include <stddef.h>
#include "cfcs.h"
#include "collision_induced_absorption.h"
#include "molecule_table.h"
#include "tables.h"

typedef struct Mol
{
    int id;
    char const *name;
    float amu;
    int num_iso;
} Mol_t;

static Mol_t const mols[] = {
#define X(id, name, amu, niso, q296, beta) {id, name, amu, niso},
    GRT_MOLECULE_TABLE(X)
#undef X
};

int grt_molecule_info(int id, char const **name, double *mass_g, int *num_iso)
{
    size_t i;
    for (i=0; i<sizeof(mols)/sizeof(mols[0]); ++i)
    {
        if (mols[i].id == id)
        {
            if (name != NULL) *name = mols[i].name;
            if (mass_g != NULL)
            {
                double m = mols[i].amu; /* float literal widened, then divided in double */
                m /= 6.023e23;
                *mass_g = m;
            }
            if (num_iso != NULL) *num_iso = mols[i].num_iso;
            return 1;
        }
    }
    return 0;
}

char const *grt_cfc_name(int id)
{
    static char const * const names[NUM_CFCS] = {
        "CFC-11", "CFC-12", "CFC-113", "CFC-114", "CFC-115", "HCFC-22", "HCFC-141b",
        "HCFC-142b", "HFC-23", "HFC-125", "HFC-134a", "HFC-143a", "HFC-152a", "HFC-227ea",
        "HFC-245fa", "CCl4", "C2F6", "CF4", "CH2Cl2", "NF3", "SF6"};
    return (id >= 0 && id < NUM_CFCS) ? names[id] : NULL;
}

char const *grt_cia_name(int id)
{
    switch (id)
    {
        case CIA_N2: return "N2";
        case CIA_O2: return "O2";
        default: return NULL;
    }
}
