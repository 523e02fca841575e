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
 * (extern/NrrdIO/CMakeLists.txt:67-69).  A header NrrdIO refuses is refused here as well (see
 * "header rules" below): the two readers agree on every file of the differential tests.
 */
#include "nrrd.h"

#include <cmath>
#include <algorithm>
#include <cctype>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <new>
#include <stdexcept>
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

/* ---- header rules ---------------------------------------------------------------------------------
 * The reference hands the file to NrrdIO (nrrd_file_parser.cpp:23-26), which refuses more than a
 * malformed size line: a header line that is neither a comment, a key:=value pair nor "field: value" with
 * a known field, a field given twice, per-axis fields before `dimension` or with another count than
 * `dimension`, space vectors before `space`, names it does not know (types, encodings, endian, centers,
 * kinds, spaces), a kind that does not fit its axis size, a missing `encoding`.  The checks below make this
 * reader refuse the same headers (tests/test_host.py compares both on a list of such files and on random
 * headers).  Values of fields the renderer does not use are checked for form only. */
struct FieldDef {
    const char *name;
    int id;
};
enum {
    F_CONTENT, F_NUMBER, F_TYPE, F_BLOCKSIZE, F_DIMENSION, F_SPACE, F_SPACEDIM, F_SIZES, F_SPACINGS, F_THICKNESSES, F_AXISMINS,
    F_AXISMAXS, F_SPACEDIRS, F_CENTERS, F_KINDS, F_LABELS, F_UNITS, F_MIN, F_MAX, F_OLDMIN, F_OLDMAX, F_ENDIAN, F_ENCODING,
    F_LINESKIP, F_BYTESKIP, F_SAMPLEUNITS, F_SPACEUNITS, F_SPACEORIGIN, F_MEASFRAME, F_DATAFILE, F_COUNT
};
const FieldDef kFields[] = {
    {"content", F_CONTENT}, {"number", F_NUMBER}, {"type", F_TYPE}, {"block size", F_BLOCKSIZE}, {"blocksize", F_BLOCKSIZE},
    {"dimension", F_DIMENSION}, {"space", F_SPACE}, {"space dimension", F_SPACEDIM}, {"sizes", F_SIZES}, {"spacings", F_SPACINGS},
    {"thicknesses", F_THICKNESSES}, {"axis mins", F_AXISMINS}, {"axismins", F_AXISMINS}, {"axis maxs", F_AXISMAXS},
    {"axismaxs", F_AXISMAXS}, {"space directions", F_SPACEDIRS}, {"centers", F_CENTERS}, {"centerings", F_CENTERS},
    {"kinds", F_KINDS}, {"labels", F_LABELS}, {"units", F_UNITS}, {"min", F_MIN}, {"max", F_MAX}, {"old min", F_OLDMIN},
    {"oldmin", F_OLDMIN}, {"old max", F_OLDMAX}, {"oldmax", F_OLDMAX}, {"endian", F_ENDIAN}, {"encoding", F_ENCODING},
    {"line skip", F_LINESKIP}, {"lineskip", F_LINESKIP}, {"byte skip", F_BYTESKIP}, {"byteskip", F_BYTESKIP},
    {"sample units", F_SAMPLEUNITS}, {"sampleunits", F_SAMPLEUNITS}, {"space units", F_SPACEUNITS},
    {"space origin", F_SPACEORIGIN}, {"measurement frame", F_MEASFRAME}, {"data file", F_DATAFILE}, {"datafile", F_DATAFILE}};

std::vector<std::string> words(const std::string &s)
{
    std::vector<std::string> w;
    std::istringstream ss(s);
    std::string t;
    while (ss >> t) w.push_back(t);
    return w;
}

bool all_doubles(const std::vector<std::string> &w)
{
    for (const std::string &t : w) {
        const std::string l = lower(t);
        if (l == "nan" || l == "inf" || l == "+inf" || l == "-inf") continue;
        char *end = nullptr;
        std::strtod(t.c_str(), &end);
        if (end == t.c_str()) return false;
    }
    return true;
}

/* "(a,b,c)" with `comps` numbers, or "none" where allowed */
bool is_vector(const std::string &t, int comps, bool none_ok)
{
    if (none_ok && lower(t) == "none") return true;
    if (t.size() < 3 || t.front() != '(' || t.back() != ')') return false;
    std::string body = t.substr(1, t.size() - 2);
    int n = 0;
    size_t a = 0;
    while (true) {
        const size_t b = body.find(',', a);
        const std::string piece = body.substr(a, b == std::string::npos ? std::string::npos : b - a);
        if (!all_doubles({trim(piece)}) || trim(piece).empty()) return false;
        n++;
        if (b == std::string::npos) break;
        a = b + 1;
    }
    return n == comps;
}

/* number of "quoted strings" in s, or -1 if it is not a sequence of them; the strings go to *out */
int quoted_count(const std::string &s, std::vector<std::string> *out = nullptr)
{
    int n = 0;
    size_t i = 0;
    while (true) {
        while (i < s.size() && std::isspace((unsigned char)s[i])) i++;
        if (i == s.size()) return n;
        if (s[i] != '"') return -1;
        const size_t b = ++i;
        while (i < s.size() && s[i] != '"') i += (s[i] == '\\' && i + 1 < s.size()) ? 2 : 1;
        if (i >= s.size()) return -1;
        if (out) out->push_back(s.substr(b, i - b));
        i++;
        n++;
    }
}

int space_dimension_of(const std::string &name)
{
    static const struct {
        const char *n;
        int d;
    } tab[] = {{"right-anterior-superior", 3}, {"ras", 3}, {"left-anterior-superior", 3}, {"las", 3}, {"left-posterior-superior", 3},
               {"lps", 3}, {"right-anterior-superior-time", 4}, {"rast", 4}, {"left-anterior-superior-time", 4}, {"last", 4},
               {"left-posterior-superior-time", 4}, {"lpst", 4}, {"scanner-xyz", 3}, {"scanner-xyz-time", 4}, {"3d-right-handed", 3},
               {"3d-left-handed", 3}, {"3d-right-handed-time", 4}, {"3d-left-handed-time", 4}};
    for (const auto &e : tab)
        if (name == e.n) return e.d;
    return 0;
}

/* size a kind demands of its axis: 0 = any, -1 = unknown kind */
int kind_size(const std::string &k)
{
    static const struct {
        const char *n;
        int s;
    } tab[] = {{"domain", 0}, {"space", 0}, {"time", 0}, {"list", 0}, {"point", 0}, {"vector", 0}, {"covariant-vector", 0},
               {"normal", 0}, {"???", 0}, {"none", 0}, {"stub", 1}, {"scalar", 1}, {"complex", 2}, {"2-vector", 2}, {"3-color", 3},
               {"rgb-color", 3}, {"hsv-color", 3}, {"xyz-color", 3}, {"4-color", 4}, {"rgba-color", 4}, {"3-vector", 3},
               {"3-gradient", 3}, {"3-normal", 3}, {"4-vector", 4}, {"quaternion", 4}, {"2d-symmetric-matrix", 3},
               {"2d-masked-symmetric-matrix", 4}, {"2d-matrix", 4}, {"2d-masked-matrix", 5}, {"3d-symmetric-matrix", 6},
               {"3d-masked-symmetric-matrix", 7}, {"3d-matrix", 9}, {"3d-masked-matrix", 10}};
    for (const auto &e : tab)
        if (k == e.n) return e.s;
    return -1;
}

template <typename T>
void store_as(uint8_t *dst, size_t i, double v, long long iv, unsigned long long uv, bool is_float, bool is_unsigned)
{
    T t = is_float ? static_cast<T>(v) : (is_unsigned ? static_cast<T>(uv) : static_cast<T>(iv));
    std::memcpy(dst + i * sizeof(T), &t, sizeof(T));
}
}  // namespace

static int nrrd_read_impl(const char *path, VrNrrd &out, std::string &err, VrNrrdPayload *payload);

/* no exception crosses the ABI: the extern "C" loaders (api.cu) call only this */
int vr_nrrd_read(const char *path, VrNrrd &out, std::string &err, VrNrrdPayload *payload)
{
    try {
        return nrrd_read_impl(path, out, err, payload);
    } catch (const std::bad_alloc &) {
        err = "Failed to read file: out of memory";
        return VR_ERR_NOMEM;
    } catch (const std::exception &e) {
        err = std::string("Failed to read file: ") + e.what();
        return VR_ERR_FORMAT;
    }
}

static int nrrd_read_impl(const char *path, VrNrrd &out, std::string &err, VrNrrdPayload *payload)
{
    std::ifstream f(path, std::ios::binary);
    if (!f) {
        err = std::string("Failed to read file: ") + path;
        return VR_ERR_IO;
    }
    std::string line;
    auto bad = [&](const std::string &why) {
        err = "Failed to read file: " + why;
        return VR_ERR_FORMAT;
    };
    if (!std::getline(f, line)) return bad("not a NRRD header");
    if (!line.empty() && line.back() == '\r') line.pop_back();
    if (line.size() != 8 || line.compare(0, 7, "NRRD000") != 0 || line[7] < '1' || line[7] > '5') return bad("not a NRRD header");
    Scalar type = BAD;
    int dimension = -1, space_dim = 0;
    std::vector<uint64_t> sizes;
    std::string encoding, endian, datafile;
    long long byte_skip = 0, line_skip = 0;
    bool seen[F_COUNT] = {};
    std::vector<std::string> spacings, axis_mins, axis_maxs, units, space_dirs, kinds;
    while (std::getline(f, line)) {
        if (!line.empty() && line.back() == '\r') line.pop_back();
        if (line.empty()) break; /* blank line: data follows (attached header) */
        if (line[0] == '#') continue;
        const size_t sep = line.find(": "), kv = line.find(":=");
        if (kv != std::string::npos && (sep == std::string::npos || kv < sep)) continue; /* key:=value pair */
        if (sep == std::string::npos) return bad("unrecognised header line \"" + line + "\"");
        const std::string key = lower(line.substr(0, sep));
        int id = -1;
        for (const FieldDef &fd : kFields)
            if (key == fd.name) id = fd.id;
        if (id < 0) return bad("unknown field \"" + line.substr(0, sep) + "\"");
        if (seen[id]) return bad("field \"" + key + "\" given twice");
        seen[id] = true;
        size_t vb = sep + 2;
        while (vb < line.size() && line[vb] == ' ') vb++;
        const std::string val = line.substr(vb);
        const bool per_axis = id == F_SIZES || id == F_SPACINGS || id == F_THICKNESSES || id == F_AXISMINS || id == F_AXISMAXS ||
                              id == F_SPACEDIRS || id == F_CENTERS || id == F_KINDS || id == F_LABELS || id == F_UNITS;
        if (per_axis && dimension < 0) return bad("\"" + key + "\" before dimension");
        const bool needs_space = id == F_SPACEDIRS || id == F_SPACEORIGIN || id == F_SPACEUNITS || id == F_MEASFRAME;
        if (needs_space && space_dim == 0) return bad("\"" + key + "\" before space");
        const std::vector<std::string> w = words(val);
        switch (id) {
            case F_TYPE:
                type = parse_type(lower(val));
                if (type == BAD) return bad("missing or unknown type");
                break;
            case F_DIMENSION:
                dimension = w.empty() ? 0 : std::atoi(w[0].c_str());
                if (dimension < 1 || dimension > 16) {
                    err = "Invalid file properties";
                    return VR_ERR_FORMAT;
                }
                break;
            case F_SIZES:
                for (const std::string &t : w) {
                    if (t.empty() || !std::isdigit((unsigned char)t[0])) return bad("bad sizes");
                    sizes.push_back(std::strtoull(t.c_str(), nullptr, 10));
                }
                if ((int)sizes.size() != dimension) {
                    err = "Invalid file properties";
                    return VR_ERR_FORMAT;
                }
                break;
            case F_SPACINGS:
                spacings = w;
                /* fall through */
            case F_THICKNESSES:
            case F_AXISMINS:
            case F_AXISMAXS:
                if ((int)w.size() != dimension || !all_doubles(w)) return bad("bad \"" + key + "\"");
                if (id == F_AXISMINS) axis_mins = w;
                if (id == F_AXISMAXS) axis_maxs = w;
                break;
            case F_CENTERS:
                if ((int)w.size() != dimension) return bad("bad centers");
                for (const std::string &t : w) {
                    const std::string l = lower(t);
                    if (l != "cell" && l != "node" && l != "???" && l != "none") return bad("bad centers");
                }
                break;
            case F_KINDS:
                if ((int)w.size() != dimension) return bad("bad kinds");
                for (const std::string &t : w) {
                    if (kind_size(lower(t)) < 0) return bad("bad kinds");
                    kinds.push_back(lower(t));
                }
                break;
            case F_LABELS:
                if (quoted_count(val) != dimension) return bad("bad labels");
                break;
            case F_UNITS:
                if (quoted_count(val, &units) != dimension) return bad("bad units");
                break;
            case F_SPACE:
                if (seen[F_SPACEDIM]) return bad("both space and space dimension");
                space_dim = space_dimension_of(lower(val));
                if (space_dim == 0) return bad("unknown space");
                break;
            case F_SPACEDIM:
                if (seen[F_SPACE]) return bad("both space and space dimension");
                space_dim = w.empty() ? 0 : std::atoi(w[0].c_str());
                if (space_dim < 1 || space_dim > 8) return bad("bad space dimension");
                break;
            case F_SPACEDIRS:
                if ((int)w.size() != dimension) return bad("bad space directions");
                for (const std::string &t : w)
                    if (!is_vector(t, space_dim, true)) return bad("bad space directions");
                space_dirs = w;
                break;
            case F_SPACEORIGIN:
                if (w.size() != 1 || !is_vector(w[0], space_dim, false)) return bad("bad space origin");
                break;
            case F_MEASFRAME:
                if ((int)w.size() != space_dim) return bad("bad measurement frame");
                for (const std::string &t : w)
                    if (!is_vector(t, space_dim, true)) return bad("bad measurement frame");
                break;
            case F_SPACEUNITS:
                if (quoted_count(val) != space_dim) return bad("bad space units");
                break;
            case F_BLOCKSIZE: return bad("block size without block type");
            case F_ENCODING: encoding = lower(val); break;
            case F_ENDIAN:
                endian = lower(val);
                if (endian != "little" && endian != "big") return bad("bad endian");
                break;
            case F_DATAFILE: datafile = trim(val); break;
            case F_BYTESKIP:
                byte_skip = std::atoll(val.c_str());
                if (byte_skip < -1) return bad("bad byte skip");
                break;
            case F_LINESKIP:
                line_skip = std::atoll(val.c_str());
                if (line_skip < 0) return bad("bad line skip");
                break;
            case F_OLDMIN:
            case F_OLDMAX:
                if (w.empty() || !all_doubles({w[0]})) return bad("bad \"" + key + "\"");
                break;
            default: break; /* content, number, min, max, sample units: free text */
        }
    }
    if (type == BAD) return bad("missing or unknown type");
    if (encoding.empty()) return bad("missing encoding");
    if (dimension != 3 || sizes.size() != 3) {
        err = "Invalid file properties"; /* nrrd_file_parser.cpp:28-30 */
        return VR_ERR_FORMAT;
    }
    for (int i = 0; i < 3; i++)
        if (sizes[i] == 0 || sizes[i] > 0xffffffffull) {
            err = "Invalid file properties";
            return VR_ERR_FORMAT;
        }
    for (size_t i = 0; i < kinds.size() && i < 3; i++) {
        const int ks = kind_size(kinds[i]);
        if (ks > 0 && (uint64_t)ks != sizes[i]) return bad("kind does not fit the axis size");
    }
    /* an axis with a space direction has no spacing, no axis min / max and no units of its own */
    for (size_t i = 0; i < space_dirs.size(); i++) {
        if (lower(space_dirs[i]) == "none") continue;
        if ((i < spacings.size() && lower(spacings[i]) != "nan") || (i < axis_mins.size() && lower(axis_mins[i]) != "nan") ||
            (i < axis_maxs.size() && lower(axis_maxs[i]) != "nan") || (i < units.size() && !units[i].empty()))
            return bad("axis with a space direction and a spacing, axis min / max or units");
    }
    for (int i = 0; i < 3; i++) {
        out.spacing[i] = __builtin_nan("");
        if ((size_t)i < space_dirs.size() && lower(space_dirs[i]) != "none") {
            /* "(a,b,c)": the length of the direction vector */
            const std::string body = space_dirs[i].substr(1, space_dirs[i].size() - 2);
            double n2 = 0.0;
            size_t a = 0;
            while (true) {
                const size_t b = body.find(',', a);
                const double v = std::strtod(trim(body.substr(a, b == std::string::npos ? std::string::npos : b - a)).c_str(), nullptr);
                n2 += v * v;
                if (b == std::string::npos) break;
                a = b + 1;
            }
            out.spacing[i] = std::sqrt(n2);
        } else if ((size_t)i < spacings.size() && lower(spacings[i]) != "nan") {
            out.spacing[i] = std::fabs(std::strtod(spacings[i].c_str(), nullptr));
        }
    }
    const size_t esz = scalar_size(type);
    /* sizes come from the header alone: refuse what cannot be a volume before anything is allocated from them
     * (each may be up to 2^32-1; the product must not wrap, and the reference could not hold it either) */
    unsigned __int128 n128 = (unsigned __int128)sizes[0] * (unsigned __int128)sizes[1] * (unsigned __int128)sizes[2];
    if (n128 == 0 || n128 * esz > ((unsigned __int128)1 << 46)) {
        err = "Failed to read file: sizes out of range";
        return VR_ERR_FORMAT;
    }
    const size_t n = (size_t)n128;
    const bool enc_ascii = encoding == "ascii" || encoding == "text" || encoding == "txt";
    if (encoding != "raw" && encoding != "hex" && !enc_ascii) return bad("unsupported encoding " + encoding);
    if (esz > 1 && !enc_ascii && endian.empty()) return bad("endian required");

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

    {
        /* raw and hex payloads cannot be shorter than the data they claim; an ascii value takes at least two bytes */
        const std::streamoff here = in->tellg();
        in->seekg(0, std::ios::end);
        const std::streamoff end = in->tellg();
        in->seekg(here, std::ios::beg);
        const unsigned __int128 need = encoding == "raw" ? (unsigned __int128)n * esz : (enc_ascii ? (unsigned __int128)n * 2 - 1 : (unsigned __int128)n * esz * 2);
        if (here < 0 || end < here || (unsigned __int128)(end - here) < need) {
            err = "Failed to read file: truncated data";
            return VR_ERR_IO;
        }
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
    } else if (enc_ascii) {
        if (byte_skip > 0) in->seekg(byte_skip, std::ios::cur);
        const bool is_float = type == F32 || type == F64, is_unsigned = type == U8 || type == U16 || type == U32 || type == U64;
        std::string tok;
        for (size_t i = 0; i < n; i++) {
            if (!(*in >> tok)) {
                err = "Failed to read file: truncated data";
                return VR_ERR_IO;
            }
            /* a value is whatever number the token starts with (sscanf); a token that starts with none is an error */
            char *e1 = nullptr, *e2 = nullptr, *e3 = nullptr;
            const double dv = is_float ? std::strtod(tok.c_str(), &e1) : 0.0;
            const long long iv = (!is_float && !is_unsigned) ? std::strtoll(tok.c_str(), &e2, 10) : 0;
            const unsigned long long uv = is_unsigned ? std::strtoull(tok.c_str(), &e3, 10) : 0;
            if ((is_float ? e1 : (is_unsigned ? e3 : e2)) == tok.c_str()) return bad("bad value \"" + tok + "\" in ascii data");
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
