/* Working precision of the library (reference: utilities/src/floating_point_type.h:23-29).
 * Default double; -DSINGLE_PRECISION (the reference's --enable-single-precision,
 * configure.ac:25-32) makes fp_t float in every public signature and struct: that build is
 * libgrtcode_sp.so + libgrt_sm100_sp.so. */
#ifndef FLOATING_POINT_TYPE_H_
#define FLOATING_POINT_TYPE_H_
#ifdef SINGLE_PRECISION
typedef float fp_t;
#define epsilon_ 1.e-7
#else
typedef double fp_t;
#define epsilon_ 1.e-12
#endif
#endif
