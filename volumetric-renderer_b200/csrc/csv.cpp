/*
 * csv.cpp -- stack-of-CSV-slices importer (host side, no GPU).
 *
 * Produces what the reference's CsvFileParser::parse produces (src/data/csv_file_parser.cpp:14-50):
 * one file per z slice in the order given, one text line per y row, comma-separated values along x.
 * Each file is read in one piece and scanned in place.  The rules that matter for equality:
 *   - pieces: a line ends at '\n', a field at ','; text after the last separator counts only if it
 *     is not empty (what std::getline does, :22/:25), so "1,2," has two fields and "1,,2" three;
 *   - every field goes through std::stof (:26): leading white space skipped, trailing characters
 *     ignored ("7\r" is 7), an empty or non-numeric field or an out-of-range value is an error
 *     reported with std::stof's own message;
 *   - the x extent is the field count of the first line of the first file, the y extent the line
 *     count of the first file; any other count is "Inconsistant dimensions" (:32-45, the
 *     reference's spelling is kept in the message);
 *   - min and max start at 0 because the reference value-initialises its Dataset (:16), so
 *     min <= 0 <= max always; NaN fields never replace either (:27-28);
 *   - a file that cannot be opened reads as zero lines (the reference never checks the stream).
 * Values are kept as float, the reference's only voxel type (src/data/dataset.h:9-13).
 */
#include <cstdint>
#include <cstdio>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/vr.h"
#include "nrrd.h"

namespace
{
/* whole file as bytes; an unreadable file is an empty one */
std::string slurp(const char *path)
{
    std::string s;
    if (FILE *f = std::fopen(path, "rb")) {
        char buf[1 << 16];
        size_t n;
        while ((n = std::fread(buf, 1, sizeof buf, f)) > 0) s.append(buf, n);
        std::fclose(f);
    }
    return s;
}

/* call fn(begin, end) for every piece of [b, e) separated by sep; a trailing empty piece is dropped */
template <typename F>
void pieces(const char *b, const char *e, char sep, F fn)
{
    while (b < e) {
        const char *q = b;
        while (q < e && *q != sep) q++;
        fn(b, q);
        b = q + 1; /* past the separator (or past the end) */
    }
}
}  // namespace

int vr_csv_read(const char *const *paths, uint32_t npaths, VrCsv &out, std::string &err)
{
    out.data.clear();
    out.dims[0] = out.dims[1] = out.dims[2] = 0;
    out.min = 0.0f, out.max = 0.0f;
    bool bad_dims = false;
    std::string field;
    try {
        for (uint32_t z = 0; z < npaths; z++) {
            if (!paths[z]) {
                err = "vr_csv_read: null path";
                return VR_ERR_INVALID;
            }
            const std::string text = slurp(paths[z]);
            uint32_t rows = 0;
            pieces(text.data(), text.data() + text.size(), '\n', [&](const char *lb, const char *le) {
                if (bad_dims) return;
                uint32_t cols = 0;
                pieces(lb, le, ',', [&](const char *fb, const char *fe) {
                    field.assign(fb, fe);
                    const float v = std::stof(field);
                    if (v < out.min) out.min = v;
                    if (out.max < v) out.max = v;
                    out.data.push_back(v);
                    cols++;
                });
                if (z == 0 && rows == 0) out.dims[0] = cols;
                else if (cols != out.dims[0]) bad_dims = true;
                rows++;
            });
            if (!bad_dims) {
                if (z == 0) out.dims[1] = rows;
                else if (rows != out.dims[1]) bad_dims = true;
            }
            if (bad_dims) {
                err = "Inconsistant dimensions";
                return VR_ERR_FORMAT;
            }
        }
    } catch (const std::exception &e) {
        err = e.what(); /* std::stof's message, as the reference's error dialog shows it */
        return VR_ERR_FORMAT;
    }
    out.dims[2] = npaths;
    return VR_OK;
}
