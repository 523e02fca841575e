/* nrrd.h -- internal interface of the native NRRD reader (nrrd.cpp). */
#ifndef VR_NRRD_H
#define VR_NRRD_H
#include <stdint.h>

#include <string>
#include <vector>

#include "../../include/vr.h"

struct VrNrrd {
    int dtype = VR_F32;          /* vr_dtype of `bytes` */
    uint32_t dims[3] = {0, 0, 0}; /* x fastest */
    /* sample distance per axis as the header states it -- `spacings`, or the length of the axis's `space directions`
     * vector -- NaN where it states none.  The reference ignores both (nrrd_file_parser.cpp:32-33) and so does an
     * upload; vr_get_volume_spacing hands them to a caller who wants them honoured. */
    double spacing[3] = {__builtin_nan(""), __builtin_nan(""), __builtin_nan("")};
    std::vector<uint8_t> bytes;
};

/* Where the payload of a file lies when it can go to the device as it is: raw encoding, a type the
 * device keeps natively (u8, u16, f32) and the host's byte order. */
struct VrNrrdPayload {
    bool direct = false;
    std::string path;    /* the file holding the data (the header itself, or its detached data file) */
    uint64_t offset = 0; /* first byte of the payload in that file */
    uint64_t bytes = 0;
};

/* Returns a vr_status; on failure `err` holds the message.  With `payload` given and the file qualifying
 * (see VrNrrdPayload) the data is NOT read: out.dtype / out.dims are set, out.bytes stays empty and
 * *payload says where to stream it from (vr_load_nrrd does that through pinned memory). */
int vr_nrrd_read(const char *path, VrNrrd &out, std::string &err, VrNrrdPayload *payload = nullptr);
#endif

/* Stack of CSV slices (csv.cpp): what the reference's CsvFileParser::parse returns. */
struct VrCsv {
    uint32_t dims[3] = {0, 0, 0}; /* x = fields per line, y = lines per file, z = files */
    float min = 0.0f, max = 0.0f; /* both start at 0, like the reference's value-initialised Dataset */
    std::vector<float> data;      /* x fastest */
};
int vr_csv_read(const char *const *paths, uint32_t npaths, VrCsv &out, std::string &err);
