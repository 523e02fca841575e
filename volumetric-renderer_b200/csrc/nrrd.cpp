/*
 * nrrd.cpp -- native NRRD / NHDR reader for the loader fast path.
 *
 * Stands where NrrdFileParser::parse + NrrdIO stand in the reference
 * (src/data/nrrd_file_parser.cpp:21-77): read a 3-D scalar dataset, x fastest.  Differences by
 * design: uchar and ushort data stay in their native type for the device (the reference widens
 * everything to float, :67-77); every other scalar type is widened with static_cast<float>
 * exactly like :49-77.  As in the reference, `dimension` must be 3 (:28-30) and axis spacings /
 * space directions are ignored (:32-33).  Encodings: raw, ascii (text), hex.  gzip/bzip2 are
 * rejected -- the reference's NrrdIO build compiles them out too
 * (extern/NrrdIO/CMakeLists.txt:67-69).
 */
#include "nrrd.h"

#include <algorithm>
#include <cctype>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <sstream>

namespace
{
enum Scalar { I8, U8, I16, U16, I32, U32, I64, U64, F32, F64, BAD };

std::string trim(const std::string &s)
{
    size_t a = 0, b = s.size();
    while (a < b && std::isspace((unsigned char)s[a])) a++;
    while (b > a && std::isspace((unsigned char)s[b - 1])) b--;
    return s.substr(a, b - a);
}

std::string lower(std::string s)
{
    for (auto &c : s) c = (char)std::tolower((unsigned char)c);
    return s;
}

Scalar parse_type(const std::string &t)
{
    /* type names of the NRRD format definition */
    static const struct {
        const char *name;
        Scalar s;
    } tab[] = {{"signed char", I8},    {"int8", I8},        {"int8_t", I8},
               {"uchar", U8},          {"unsigned char", U8}, {"uint8", U8},      {"uint8_t", U8},
               {"short", I16},         {"short int", I16},  {"signed short", I16}, {"signed short int", I16},
               {"int16", I16},         {"int16_t", I16},
               {"ushort", U16},        {"unsigned short", U16}, {"unsigned short int", U16}, {"uint16", U16},
               {"uint16_t", U16},
               {"int", I32},           {"signed int", I32}, {"int32", I32},     {"int32_t", I32},
               {"uint", U32},          {"unsigned int", U32}, {"uint32", U32},  {"uint32_t", U32},
               {"longlong", I64},      {"long long", I64},  {"long long int", I64}, {"signed long long", I64},
               {"signed long long int", I64}, {"int64", I64}, {"int64_t", I64},
               {"ulonglong", U64},     {"unsigned long long", U64}, {"unsigned long long int", U64},
               {"uint64", U64},        {"uint64_t", U64},
               {"float", F32},         {"double", F64}};
    for (const auto &e : tab)
        if (t == e.name) return e.s;
    return BAD;
}

size_t scalar_size(Scalar s)
{
    switch (s) {
        case I8: case U8: return 1;
        case I16: case U16: return 2;
        case I32: case U32: case F32: return 4;
        default: return 8;
    }
}

bool host_is_little()
{
    const uint16_t one = 1;
    return *reinterpret_cast<const uint8_t *>(&one) == 1;
}

template <typename T>
void widen(const uint8_t *src, size_t n, float *dst)
{
    for (size_t i = 0; i < n; i++) {
        T v;
        std::memcpy(&v, src + i * sizeof(T), sizeof(T));
        dst[i] = static_cast<float>(v); /* nrrd_file_parser.cpp:74 */
    }
}

template <typename T>
void store_as(uint8_t *dst, size_t i, double v, long long iv, unsigned long long uv, bool is_float, bool is_unsigned)
{
    T t = is_float ? static_cast<T>(v) : (is_unsigned ? static_cast<T>(uv) : static_cast<T>(iv));
    std::memcpy(dst + i * sizeof(T), &t, sizeof(T));
}
}  // namespace

int vr_nrrd_read(const char *path, VrNrrd &out, std::string &err, VrNrrdPayload *payload)
{
    std::ifstream f(path, std::ios::binary);
    if (!f) {
        err = std::string("Failed to read file: ") + path;
        return VR_ERR_IO;
    }
    std::string line;
    if (!std::getline(f, line) || line.compare(0, 4, "NRRD") != 0) {
        err = "Failed to read file: not a NRRD header";
        return VR_ERR_FORMAT;
    }
    Scalar type = BAD;
    int dimension = -1;
    std::vector<uint64_t> sizes;
    std::string encoding = "raw", endian, datafile;
    long long byte_skip = 0, line_skip = 0;
    while (std::getline(f, line)) {
        if (!line.empty() && line.back() == '\r') line.pop_back();
        if (line.empty()) break; /* blank line: data follows (attached header) */
        if (line[0] == '#') continue;
        const size_t colon = line.find(':');
        if (colon == std::string::npos) continue;
        if (colon + 1 < line.size() && line[colon + 1] == '=') continue; /* key:=value pair */
        const std::string key = lower(trim(line.substr(0, colon)));
        const std::string val = trim(line.substr(colon + 1));
        if (key == "type") type = parse_type(lower(val));
        else if (key == "dimension") dimension = std::atoi(val.c_str());
        else if (key == "sizes") {
            std::istringstream ss(val);
            uint64_t v;
            while (ss >> v) sizes.push_back(v);
        } else if (key == "encoding") encoding = lower(val);
        else if (key == "endian") endian = lower(val);
        else if (key == "data file" || key == "datafile") datafile = val;
        else if (key == "byte skip" || key == "byteskip") byte_skip = std::atoll(val.c_str());
        else if (key == "line skip" || key == "lineskip") line_skip = std::atoll(val.c_str());
    }
    if (type == BAD) {
        err = "Failed to read file: missing or unknown type";
        return VR_ERR_FORMAT;
    }
    if (dimension != 3 || sizes.size() != 3) {
        err = "Invalid file properties"; /* nrrd_file_parser.cpp:28-30 */
        return VR_ERR_FORMAT;
    }
    for (int i = 0; i < 3; i++)
        if (sizes[i] == 0 || sizes[i] > 0xffffffffull) {
            err = "Invalid file properties";
            return VR_ERR_FORMAT;
        }
    const size_t n = (size_t)sizes[0] * sizes[1] * sizes[2];
    const size_t esz = scalar_size(type);

    /* where the payload is */
    std::ifstream df;
    std::istream *in = &f;
    std::string payload_path(path);
    if (!datafile.empty()) {
        if (datafile.find("LIST") == 0 || datafile.find('%') != std::string::npos) {
            err = "Failed to read file: multi-file data not supported";
            return VR_ERR_FORMAT;
        }
        std::string p = datafile;
        if (p[0] != '/') {
            const std::string hp(path);
            const size_t slash = hp.find_last_of('/');
            if (slash != std::string::npos) p = hp.substr(0, slash + 1) + p;
        }
        df.open(p, std::ios::binary);
        if (!df) {
            err = "Failed to read file: data file " + p;
            return VR_ERR_IO;
        }
        in = &df;
        payload_path = p;
    }
    for (long long i = 0; i < line_skip; i++) std::getline(*in, line);

    if (payload && encoding == "raw" && (type == U8 || type == U16 || type == F32) && byte_skip >= -1 &&
        (esz == 1 || (!endian.empty() && (endian == "little") == host_is_little()))) {
        const std::streamoff here = in->tellg();
        in->seekg(0, std::ios::end);
        const std::streamoff end = in->tellg();
        const std::streamoff off = byte_skip == -1 ? end - (std::streamoff)(n * esz) : here + (std::streamoff)byte_skip;
        if (here < 0 || end < 0 || off < 0 || off + (std::streamoff)(n * esz) > end) {
            err = "Failed to read file: truncated data";
            return VR_ERR_IO;
        }
        payload->direct = true, payload->path = payload_path, payload->offset = (uint64_t)off, payload->bytes = n * esz;
        out.dims[0] = (uint32_t)sizes[0], out.dims[1] = (uint32_t)sizes[1], out.dims[2] = (uint32_t)sizes[2];
        out.dtype = type == U8 ? VR_U8 : (type == U16 ? VR_U16 : VR_F32);
        out.bytes.clear();
        return VR_OK;
    }

    std::vector<uint8_t> rawbytes(n * esz);
    if (encoding == "raw") {
        if (byte_skip == -1) {
            in->seekg(-(std::streamoff)(n * esz), std::ios::end);
        } else if (byte_skip > 0) {
            in->seekg(byte_skip, std::ios::cur);
        }
        in->read(reinterpret_cast<char *>(rawbytes.data()), (std::streamsize)(n * esz));
        if ((size_t)in->gcount() != n * esz) {
            err = "Failed to read file: truncated data";
            return VR_ERR_IO;
        }
        if (esz > 1) {
            if (endian.empty()) {
                err = "Failed to read file: endian required";
                return VR_ERR_FORMAT;
            }
            const bool file_little = endian == "little";
            if (file_little != host_is_little())
                for (size_t i = 0; i < n; i++) std::reverse(rawbytes.begin() + i * esz, rawbytes.begin() + (i + 1) * esz);
        }
    } else if (encoding == "ascii" || encoding == "text" || encoding == "txt") {
        if (byte_skip > 0) in->seekg(byte_skip, std::ios::cur);
        const bool is_float = type == F32 || type == F64, is_unsigned = type == U8 || type == U16 || type == U32 || type == U64;
        std::string tok;
        for (size_t i = 0; i < n; i++) {
            if (!(*in >> tok)) {
                err = "Failed to read file: truncated data";
                return VR_ERR_IO;
            }
            const double dv = is_float ? std::strtod(tok.c_str(), nullptr) : 0.0;
            const long long iv = (!is_float && !is_unsigned) ? std::strtoll(tok.c_str(), nullptr, 10) : 0;
            const unsigned long long uv = is_unsigned ? std::strtoull(tok.c_str(), nullptr, 10) : 0;
            uint8_t *d = rawbytes.data();
            switch (type) {
                case I8: store_as<int8_t>(d, i, dv, iv, uv, is_float, is_unsigned); break;
                case U8: store_as<uint8_t>(d, i, dv, iv, uv, is_float, is_unsigned); break;
                case I16: store_as<int16_t>(d, i, dv, iv, uv, is_float, is_unsigned); break;
                case U16: store_as<uint16_t>(d, i, dv, iv, uv, is_float, is_unsigned); break;
                case I32: store_as<int32_t>(d, i, dv, iv, uv, is_float, is_unsigned); break;
                case U32: store_as<uint32_t>(d, i, dv, iv, uv, is_float, is_unsigned); break;
                case I64: store_as<int64_t>(d, i, dv, iv, uv, is_float, is_unsigned); break;
                case U64: store_as<uint64_t>(d, i, dv, iv, uv, is_float, is_unsigned); break;
                case F32: store_as<float>(d, i, dv, iv, uv, is_float, is_unsigned); break;
                default: store_as<double>(d, i, dv, iv, uv, is_float, is_unsigned); break;
            }
        }
    } else if (encoding == "hex") {
        if (byte_skip > 0) in->seekg(byte_skip, std::ios::cur);
        /* two hex digits per byte, most significant byte first, whitespace ignored */
        size_t got = 0;
        int hi = -1, ch;
        std::vector<uint8_t> be(n * esz);
        while (got < n * esz && (ch = in->get()) != EOF) {
            int v;
            if (ch >= '0' && ch <= '9') v = ch - '0';
            else if (ch >= 'a' && ch <= 'f') v = ch - 'a' + 10;
            else if (ch >= 'A' && ch <= 'F') v = ch - 'A' + 10;
            else continue;
            if (hi < 0) hi = v;
            else be[got++] = (uint8_t)(hi * 16 + v), hi = -1;
        }
        if (got != n * esz) {
            err = "Failed to read file: truncated data";
            return VR_ERR_IO;
        }
        rawbytes = be;
        if (esz > 1) {
            const bool file_little = endian == "little"; /* hex bytes follow the endian field like raw */
            if (file_little != host_is_little())
                for (size_t i = 0; i < n; i++) std::reverse(rawbytes.begin() + i * esz, rawbytes.begin() + (i + 1) * esz);
        }
    } else {
        err = "Failed to read file: unsupported encoding " + encoding;
        return VR_ERR_FORMAT;
    }

    out.dims[0] = (uint32_t)sizes[0], out.dims[1] = (uint32_t)sizes[1], out.dims[2] = (uint32_t)sizes[2];
    if (type == U8 || type == U16) {
        out.dtype = type == U8 ? VR_U8 : VR_U16;
        out.bytes.swap(rawbytes);
        return VR_OK;
    }
    out.dtype = VR_F32;
    out.bytes.resize(n * sizeof(float));
    float *dst = reinterpret_cast<float *>(out.bytes.data());
    switch (type) {
        case I8: widen<int8_t>(rawbytes.data(), n, dst); break;
        case I16: widen<int16_t>(rawbytes.data(), n, dst); break;
        case I32: widen<int32_t>(rawbytes.data(), n, dst); break;
        case U32: widen<uint32_t>(rawbytes.data(), n, dst); break;
        case I64: widen<int64_t>(rawbytes.data(), n, dst); break;
        case U64: widen<uint64_t>(rawbytes.data(), n, dst); break;
        case F32: widen<float>(rawbytes.data(), n, dst); break;
        default: widen<double>(rawbytes.data(), n, dst); break;
    }
    return VR_OK;
}
