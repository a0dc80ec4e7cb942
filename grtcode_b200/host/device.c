/* GPU discovery (reference: utilities/src/device.c:26-75). */
#include "device.h"
#include "errors.h"
#include "grt_sm100.h"

int get_num_gpus(int * num_devices, int const verbose)
{
    REQUIRE_PTR(num_devices);
    if (grt_device_count(num_devices) != 0)
    {
        *num_devices = 0;
    }
    if (verbose)
    {
        LOG_MESG("Found %d GPU devices:", *num_devices);
        int i;
        for (i=0; i<*num_devices; ++i)
        {
            char name[256];
            if (grt_device_name(i, name, sizeof(name)) == 0)
            {
                LOG_MESG("\tDevice #%d: %s", i, name);
            }
        }
    }
    return GRTCODE_SUCCESS;
}

int create_device(Device_t * const device, int const * const id)
{
    REQUIRE_PTR(device);
    int n;
    TRY(get_num_gpus(&n, 0));
    if (n < 1)
    {
        FAIL(GRTCODE_GPU_ERR, "no CUDA device is visible and this library has no host code path.");
    }
    int const want = id != NULL ? *id : DEFAULT_GPU;
    REQUIRE_GPU(want);
    REQUIRE_RANGE(want, 0, n - 1);
    *device = want;
    return GRTCODE_SUCCESS;
}
