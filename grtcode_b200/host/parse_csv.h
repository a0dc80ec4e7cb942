/* CSV table reader (reference: utilities/src/parse_csv.h:27-33). */
#ifndef GRT_PARSE_CSV_H_
#define GRT_PARSE_CSV_H_
#include "floating_point_type.h"

/* Reads a numeric CSV with one header line into column-major doubles:
 * values[col*num_rows + row].  Caller frees *values with free(). */
int grt_read_csv(char const *path, int *num_rows, int *num_cols, fp_t **values);

/* Reads an (x, y, ...) table, checks the column count, and resamples column 1
 * onto w0 + i*dw, i < n, with the reference's recipe (zero outside the table). */
int grt_table_samples(char const *path, int expected_cols, fp_t **table, int *rows);
int grt_table_on_grid(char const *path, int expected_cols, double w0, double dw,
                      unsigned long long n, fp_t **on_grid);
#endif
