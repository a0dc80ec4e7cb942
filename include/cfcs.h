/* CFC / HCFC / HFC cross-section species (reference: gas-optics/src/cfcs.h:32-56). */
#ifndef CFCS_H_
#define CFCS_H_
#include "grtcode_utilities.h"

#define CFC_NAME_LEN 16

typedef enum CfcId
{
    CFC11 = 0, CFC12, CFC113, CFC114, CFC115,
    HCFC22, HCFC141b, HCFC142b,
    HFC23, HFC125, HFC134a, HFC143a, HFC152a, HFC227ea, HFC245fa,
    CCl4, C2F6, CF4, CH2Cl2, NF3, SF6,
    NUM_CFCS
} CfcId_t;
#endif
