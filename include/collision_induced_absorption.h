/* Collision-induced absorption species
 * (reference: gas-optics/src/collision_induced_absorption.h:30-42). */
#ifndef COLLISION_INDUCED_ABSORPTION_H_
#define COLLISION_INDUCED_ABSORPTION_H_
#include "grtcode_utilities.h"

#define CIA_NAME_LEN 8
#define MAX_NUM_CIAS 3 /* N2-N2, N2-O2, O2-O2 */

typedef enum CiaId
{
    CIA_N2 = 0,
    CIA_O2,
    NUM_CIAS
} CiaId_t;
#endif
