/* C linkage marker (reference: utilities/src/extern.h:19-26). */
#ifndef EXTERN_H_
#define EXTERN_H_
#ifdef __cplusplus
#define EXTERN extern "C"
#else
#define EXTERN extern
#endif
#endif
