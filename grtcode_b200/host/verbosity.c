/* Verbosity level and the process-global error buffer
 * (reference: utilities/src/verbosity.c:28-83; not thread-safe there either). */
#include <string.h>
#include "errors.h"
#include "grt_sm100.h"

#define BUFFER_SIZE 4096
static int level = GRTCODE_NONE;
static char buffer[BUFFER_SIZE];

void grt_error_reset(void)
{
    buffer[0] = '\0';
}

void grt_error_append(char const *text)
{
    size_t const used = strlen(buffer);
    if (used + 1 < BUFFER_SIZE)
    {
        snprintf(buffer + used, BUFFER_SIZE - used, "%s", text);
    }
}

void grt_error_frame(char const *file, int line)
{
    char s[512];
    snprintf(s, sizeof(s), "\n\t%s: %d", file, line);
    grt_error_append(s);
}

int grt_gpu_failure(char const *file, int line)
{
    char m[1024];
    grt_last_error(m, sizeof(m));
    grt_error_reset();
    grt_error_append("Error: ");
    grt_error_append(m);
    grt_error_append("\nBacktrace:");
    grt_error_frame(file, line);
    return GRTCODE_GPU_ERR;
}

int grtcode_errstr(int const code, char * const out, int const out_size)
{
    if (out == NULL || out_size < 1)
    {
        return GRTCODE_NULL_ERR;
    }
    if (code == GRTCODE_SUCCESS)
    {
        snprintf(out, out_size, "No errors.");
    }
    else
    {
        snprintf(out, out_size, "%s\n", buffer);
    }
    return GRTCODE_SUCCESS;
}

void grtcode_set_verbosity(int const new_level)
{
    level = new_level;
}

int grtcode_verbosity(void)
{
    return level;
}
