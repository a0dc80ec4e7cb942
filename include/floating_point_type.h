/* Working precision of the library (reference: utilities/src/floating_point_type.h:23-29).
 * This build is double precision; the reference's --enable-single-precision
 * variant is not provided yet. */
#ifndef FLOATING_POINT_TYPE_H_
#define FLOATING_POINT_TYPE_H_
#ifdef SINGLE_PRECISION
#error "the sm_100a build of libgrtcode is double precision only"
#endif
typedef double fp_t;
#endif
