/* Device selection (reference: utilities/src/device.h:26-48).
 * Device ids >= 0 select a CUDA GPU.  HOST_ONLY is kept as a symbol for source
 * compatibility, but every object constructor of this library rejects it with
 * GRTCODE_VALUE_ERR: there is no CPU code path. */
#ifndef DEVICE_H_
#define DEVICE_H_
#include "extern.h"

enum grtcode_device_ids
{
    HOST_ONLY = -1,
    DEFAULT_GPU = 0
};

typedef int Device_t;

/* id == NULL selects DEFAULT_GPU (the reference falls back to HOST_ONLY when no
 * GPU exists; here that situation is an error). */
EXTERN int create_device(Device_t * const device, int const * const id);
EXTERN int get_num_gpus(int * num_devices, int const verbose);
#endif
