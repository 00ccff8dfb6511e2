// fb_csv.hpp -- csv_out of the reference (csvfile.hpp:30-117; gzip as built with USE_BOOST = 1, Makevars:1-2)
#pragma once
#include <cstdint>

namespace fb {

// Writes `header` and nrows rows of ncols values ("%.4f", comma separated, '\n' per row) to `path`: one gzip member
// when `gzip`, plain text otherwise (the reference's USE_BOOST = 0 variant). Failures are reported on stdout like the
// reference does (fast_dataframe.cpp:52-55) and returned as false.
bool write_csv_gz(const char *path, const char *header, int64_t nrows, int ncols, const double *const *cols, bool gzip);

} // namespace fb
