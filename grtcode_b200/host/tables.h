/* Names and masses of the species the library knows. */
#ifndef GRT_TABLES_H_
#define GRT_TABLES_H_
/* HITRAN molecule: name, mass [g] (= amu/6.023e23, molecules.c:307) and
 * isotopologue count; returns 0 for an unknown id. */
int grt_molecule_info(int id, char const **name, double *mass_g, int *num_iso);
char const *grt_cfc_name(int id);
char const *grt_cia_name(int id);
#endif
