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
    std::vector<uint8_t> bytes;
};

/* Returns a vr_status; on failure `err` holds the message. */
int vr_nrrd_read(const char *path, VrNrrd &out, std::string &err);
#endif
