// bgr_octree_to_boxes — OcTree ingestion: the serialised Octomap the planner
// receives in octomap_Callback (MS:273-285, octomap_msgs::fullMsgToMap -> OcTree::readData) and
// the .ot / .bt files of the same streams, turned into the fcl::OcTree::toBoxes() rows
// {x, y, z, size, occupancy, threshold} (MS:291) that bgr_map_set_boxes / bgr_map_refresh_boxes
// rasterise.  Host code: the stream is a depth-first recursion whose node sizes depend on the
// bytes already read, so it is parsed serially; the rasterisation is the device part (map.cu).
//
// Stream grammar (octomap 1.8, restated from its published file format — the library itself is
// absent from this image and from /root/reference):
//   full  (OcTree::writeData / .ot data)   node := f32 log-odds [+ payload] , u8 child mask ,
//                                                   node for every set mask bit, ascending
//   binary(OcTree::writeBinaryData / .bt)  node := u8 children 0-3 , u8 children 4-7, two bits per
//                                                   child (bit 2i, bit 2i+1): 10 free leaf (value 1),
//                                                   01 occupied leaf (value 2), 11 inner (value 3),
//                                                   00 unknown; then node for every inner child
//   child i sits at +x for bit 0, +y for bit 1, +z for bit 2; depth 16, key origin 32768.
#include <cmath>
#include <cstring>
#include <string>
#include <vector>
#include "common.cuh"

using namespace bgr;

namespace {

constexpr int kTreeDepth = 16;
constexpr int64_t kKeyOrigin = 32768;

struct Reader {
    const uint8_t* p;
    const uint8_t* end;
    bool fail = false;
    bool take(void* dst, size_t n) {
        if ((size_t)(end - p) < n) { fail = true; return false; }
        std::memcpy(dst, p, n);
        p += n;
        return true;
    }
};

struct Emit {
    double res, thr;
    std::vector<double> rows;
    double lo[3] = {1e300, 1e300, 1e300}, hi[3] = {-1e300, -1e300, -1e300};
    void leaf(int depth, const int64_t k[3], double occupancy) {
        if (!(occupancy >= thr)) return;                       // fcl::OcTree::isNodeOccupied
        const int64_t cells = (int64_t)1 << (kTreeDepth - depth);
        const double size = res * (double)cells;               // OcTreeBaseImpl::getNodeSize(depth)
        double c[3];
        for (int a = 0; a < 3; ++a) {
            // keyToCoord(key, depth): (floor((key - 32768) / 2^(16-depth)) + 0.5) * nodeSize
            c[a] = (std::floor((double)(k[a] - kKeyOrigin) / (double)cells) + 0.5) * size;
            lo[a] = std::fmin(lo[a], c[a] - 0.5 * size);
            hi[a] = std::fmax(hi[a], c[a] + 0.5 * size);
        }
        rows.insert(rows.end(), {c[0], c[1], c[2], size, occupancy, thr});
    }
};

inline double probability(float logodds) { return 1. - (1. / (1. + std::exp((double)logodds))); }   // octomap_utils.h

void child_key(const int64_t k[3], int depth, int i, int64_t out[3]) {
    const int64_t half = (int64_t)1 << (kTreeDepth - depth - 1);
    out[0] = k[0] + ((i & 1) ? half : 0);
    out[1] = k[1] + ((i & 2) ? half : 0);
    out[2] = k[2] + ((i & 4) ? half : 0);
}

bool parse_full(Reader& r, Emit& e, int depth, const int64_t k[3], int payload) {
    float value;
    uint8_t skip[8], mask;
    if (!r.take(&value, 4) || (payload && !r.take(skip, (size_t)payload)) || !r.take(&mask, 1)) return false;
    if (mask == 0) {
        e.leaf(depth, k, probability(value));
        return true;
    }
    if (depth >= kTreeDepth) { r.fail = true; return false; }
    for (int i = 0; i < 8; ++i)
        if ((mask >> i) & 1) {
            int64_t ck[3];
            child_key(k, depth, i, ck);
            if (!parse_full(r, e, depth + 1, ck, payload)) return false;
        }
    return true;
}

bool parse_binary(Reader& r, Emit& e, int depth, const int64_t k[3], double occ_leaf) {
    uint8_t c[2];
    if (!r.take(c, 2)) return false;
    if (depth >= kTreeDepth) { r.fail = true; return false; }
    for (int i = 0; i < 8; ++i) {
        const int code = (c[i >> 2] >> (2 * (i & 3))) & 3;
        if (code == 0 || code == 1) continue;                   // unknown / free leaf
        int64_t ck[3];
        child_key(k, depth, i, ck);
        if (code == 2) e.leaf(depth + 1, ck, occ_leaf);
        else if (!parse_binary(r, e, depth + 1, ck, occ_leaf)) return false;
    }
    return true;
}

}  // namespace

extern "C" {

int bgr_octree_to_boxes(const void* data, int64_t nbytes, int format, double res, double occ_threshold,
                        double* boxes_Bx6, int64_t* nb, double* res_out, double bbox_min[3], double bbox_max[3]) {
    BGR_ARG(nb != nullptr && nbytes >= 0 && (nbytes == 0 || data != nullptr), "bad octree arguments");
    Reader r{static_cast<const uint8_t*>(data), static_cast<const uint8_t*>(data) + nbytes};
    int payload = 0;
    bool binary = format == BGR_OCTREE_BINARY_DATA;
    if (format == BGR_OCTREE_FILE) {
        // text header: "# Octomap OcTree [binary ]file", comment lines, "id X", "size N", "res R", "data"
        std::string id;
        bool first = true, have_res = false, in_data = false;
        while (r.p < r.end && !in_data) {
            const uint8_t* nl = static_cast<const uint8_t*>(std::memchr(r.p, '\n', (size_t)(r.end - r.p)));
            BGR_ARG(nl != nullptr, "octree file: header is not terminated");
            std::string line(reinterpret_cast<const char*>(r.p), (size_t)(nl - r.p));
            r.p = nl + 1;
            if (!line.empty() && line.back() == '\r') line.pop_back();
            if (first) {
                first = false;
                if (line.rfind("# Octomap OcTree binary file", 0) == 0) binary = true;
                else if (line.rfind("# Octomap OcTree file", 0) == 0) binary = false;
                else BGR_ARG(false, "octree file: unknown first line '%.40s'", line.c_str());
                continue;
            }
            if (line.empty() || line[0] == '#') continue;
            if (line == "data") in_data = true;
            else if (line.rfind("id ", 0) == 0) id = line.substr(3);
            else if (line.rfind("res ", 0) == 0) { res = std::strtod(line.c_str() + 4, nullptr); have_res = true; }
            else if (line.rfind("size ", 0) == 0) {}
            else BGR_ARG(false, "octree file: unknown header line '%.40s'", line.c_str());
        }
        BGR_ARG(in_data && have_res, "octree file: header lacks 'res' or 'data'");
        if (id == "ColorOcTree") payload = 3;                   // r, g, b after the log-odds
        else if (id == "OcTreeStamped") payload = 4;            // u32 timestamp
        else if (id != "OcTree") {
            set_error("octree file: tree id '%.40s' is not supported", id.c_str());
            return BGR_ERR_UNSUPPORTED;
        }
        BGR_ARG(!(binary && payload), "binary streams carry plain OcTree nodes only");
    } else {
        BGR_ARG(format == BGR_OCTREE_FULL_DATA || format == BGR_OCTREE_BINARY_DATA, "unknown octree format %d", format);
    }
    BGR_ARG(res > 0 && std::isfinite(res), "resolution must be > 0");
    Emit e;
    e.res = res;
    e.thr = occ_threshold;
    const int64_t root[3] = {0, 0, 0};
    if (r.p < r.end) {
        // binary leaves carry the clamping maximum: logodds(0.971) stored as float (OcTree defaults)
        const double occ_leaf = probability((float)std::log(0.971 / (1. - 0.971)));
        const bool ok = binary ? parse_binary(r, e, 0, root, occ_leaf) : parse_full(r, e, 0, root, payload);
        BGR_ARG(ok && !r.fail, "octree stream is truncated or deeper than %d levels", kTreeDepth);
    }
    const int64_t n = (int64_t)e.rows.size() / 6;
    if (boxes_Bx6) {
        BGR_ARG(*nb >= n, "box buffer too small (%lld < %lld)", (long long)*nb, (long long)n);
        std::memcpy(boxes_Bx6, e.rows.data(), e.rows.size() * sizeof(double));
    }
    *nb = n;
    if (res_out) *res_out = res;
    for (int a = 0; a < 3; ++a) {
        if (bbox_min) bbox_min[a] = n ? e.lo[a] : 0.0;
        if (bbox_max) bbox_max[a] = n ? e.hi[a] : 0.0;
    }
    return BGR_OK;
}

}  // extern "C"
