/* Return codes shared with the reference (utilities/src/return_codes.h:25-40):
 * same names, same numeric values, so callers can switch libraries. */
#ifndef RETURN_CODES_H_
#define RETURN_CODES_H_
enum grtcode_return_codes
{
    GRTCODE_SUCCESS = 0,
    GRTCODE_INVALID_ERR = 1,
    GRTCODE_DIVBYZERO_ERR = 2,
    GRTCODE_OVERFLOW_ERR = 3,
    GRTCODE_UNDERFLOW_ERR = 4,
    GRTCODE_SENTINEL_ERR = 5,
    GRTCODE_NULL_ERR = 6,
    GRTCODE_NON_NULL_ERR = 7,
    GRTCODE_RANGE_ERR = 8,
    GRTCODE_VALUE_ERR = 9,
    GRTCODE_COMPILER_ERR = 10,
    GRTCODE_IO_ERR = 11,
    GRTCODE_GPU_ERR = 12
};
#endif
