/* TEST INFRASTRUCTURE (oracle).  The reference's gas-optics/src/tips2017.c is
 * missing from the mount (.MISSING_LARGE_BLOBS:1); its two symbols
 * (gas-optics/src/tips2017.h:29-37) are provided here by forwarding to the
 * substitute partition sums in grtcode_b200/host/tips.c, so that the compiled
 * reference and the new library evaluate the identical function. */
#include "../grtcode_b200/host/tips.h"

int inittips_d(void)
{
    return 0;
}

#ifdef SINGLE_PRECISION
float Q(int const mol_id, float const T, int const iso)
{
    return (float)grt_tips_q(mol_id, (double)T, iso);
}
#else
double Q(int const mol_id, double const T, int const iso)
{
    return grt_tips_q(mol_id, T, iso);
}
#endif
