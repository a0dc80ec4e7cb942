/* Total internal partition sums Q(molecule, T, isotopologue).
 *
 * Replaces the reference's gas-optics/src/tips2017.c (interface:
 * gas-optics/src/tips2017.h:29-37), whose source is absent from the reference
 * mount (.MISSING_LARGE_BLOBS:1).  The values are a SUBSTITUTE, not TIPS-2017:
 * a power law per isotopologue, corrected so that every partition-sum value the
 * reference's own tests pin is reproduced (gas-optics/test/test_tips2017.c:34-65,
 * gas-optics/test/test_kernels.c:180-217).  The function is evaluated as a
 * piecewise-linear interpolation of a 1 K table so that the host (C) and the
 * device (CUDA, un-fused IEEE ops) produce bit-identical results.
 */
#ifndef GRT_TIPS_H_
#define GRT_TIPS_H_

#ifdef __cplusplus
extern "C" {
#endif

#define GRT_TIPS_TMIN 20        /* first table temperature [K] */
#define GRT_TIPS_NT 1181        /* table entries: 20 K .. 1200 K in 1 K steps */
#define GRT_TIPS_MAX_ISO 18     /* largest isotopologue count (O3, molecules.c:48) */
#define GRT_TIPS_NUM_MOLS 53

/* Q(mol, T, iso); mol = HITRAN id 1..53, iso = 1-based isotopologue index.
 * Out-of-table ids return 1; T is clamped to the table range. */
double grt_tips_q(int mol, double T, int iso);

/* Pointer to the GRT_TIPS_NT-entry table of (mol, iso) (for the device upload). */
double const *grt_tips_table(int mol, int iso);

/* 1 if the molecule's Q is the built-in substitute, 0 if an external table supplied it
 * (GRT_TIPS_TABLE=<file> with rows "mol iso T Q", read once per process; see tips.c). */
int grt_tips_is_substitute(int mol);
/* Path of the external table in use ("" if none); *rows = rows taken from it. */
char const *grt_tips_external(int *rows);

/* Number of isotopologues the molecule table lists (molecules.c:36-306). */
int grt_tips_num_iso(int mol);

#ifdef __cplusplus
}
#endif
#endif
