/* TEST INFRASTRUCTURE — CPU oracle for the ORB extractor. Never shipped, never on the product path;
 * only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may call it.
 *
 * Plain C++ restatement of cslam/src/ORBextractor.cpp (reference, read-only at /root/reference), with every
 * intermediate stage exposed so the CUDA kernels can be compared stage by stage. Each function cites the
 * reference lines it follows. OpenCV arithmetic comes from cv_primitives.cpp (pinned against cv2 4.13.0).
 *
 * PARITY PIN: the whole pipeline is checked bit-for-bit (keypoints, order, angles, descriptors) against the
 * reference's own ORBextractor.cpp compiled unmodified (oracle/_ref/libref_orb_mono.so) by
 * tests/test_oracle_vs_ref.py, and against fixtures generated from it in tests/golden/.
 *
 * Quadtree tie rule: the reference orders equal-sized nodes by heap address (ORBextractor.cpp:852 sorts
 * pair<int,ExtractorNode*>), so its own output changes with malloc history (measured: ~3 % of keypoints differ
 * between two runs on the same image in one process). The oracle fixes the rule "addresses grow with creation
 * order", i.e. among equal sizes the LATER-created node is split first — exactly what the reference does under a
 * monotone allocator (libref_orb_mono.so). */
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <vector>

#include "cv_primitives.h"

namespace {

const int PATCH_SIZE = 31;        /* ORBextractor.cpp:63 */
const int HALF_PATCH_SIZE = 15;   /* :64 */
const int EDGE_THRESHOLD = 19;    /* :65 */

struct Tap { int x1, y1, x2, y2; };
const Tap kPattern[256] = {
#include "../motioncheck_ccm_slam_b200/csrc/brief_pattern.inc"
};

struct Kp { float x, y, size, angle, response; int octave; };   /* == ccm_keypoint (include/ccm_orb.h) */

struct Corner { int x, y, score; };   /* level-image coordinates */

struct Level {
    int w, h;
    std::vector<uint8_t> img, blur;
    std::vector<Corner> corners;  /* vToDistributeKeys, reference order, level-image coords */
    std::vector<Kp> kps;          /* after DistributeOctTree + orientation, level coords (unscaled) */
};

struct Oracle {
    int nfeatures, nlevels, iniTh, minTh;
    double scaleFactor;
    std::vector<float> sf, isf, sig2, isig2;
    std::vector<int> quota, umax;
    std::vector<Level> lv;
};

/* ORBextractor::ORBextractor, ORBextractor.cpp:579-639 (float/double mix kept as written) */
void build_tables(Oracle& o, int nfeatures, float scale, int nlevels, int ini, int min) {
    o.nfeatures = nfeatures; o.nlevels = nlevels; o.iniTh = ini; o.minTh = min;
    o.scaleFactor = scale;
    o.sf.assign(nlevels, 1.0f); o.sig2.assign(nlevels, 1.0f);
    for (int i = 1; i < nlevels; i++) {
        o.sf[i] = (float)(o.sf[i - 1] * o.scaleFactor);
        o.sig2[i] = o.sf[i] * o.sf[i];
    }
    o.isf.resize(nlevels); o.isig2.resize(nlevels);
    for (int i = 0; i < nlevels; i++) { o.isf[i] = 1.0f / o.sf[i]; o.isig2[i] = 1.0f / o.sig2[i]; }
    o.quota.assign(nlevels, 0);
    float factor = (float)(1.0f / o.scaleFactor);
    float nDesired = nfeatures * (1 - factor) / (1 - (float)std::pow((double)factor, (double)nlevels));
    int sum = 0;
    for (int l = 0; l < nlevels - 1; l++) {
        o.quota[l] = ocv_round_f(nDesired);
        sum += o.quota[l];
        nDesired *= factor;
    }
    o.quota[nlevels - 1] = std::max(nfeatures - sum, 0);
    /* umax, :623-638 */
    o.umax.assign(HALF_PATCH_SIZE + 1, 0);
    int v, v0, vmax = (int)std::floor(HALF_PATCH_SIZE * std::sqrt(2.f) / 2 + 1);
    int vmin = (int)std::ceil(HALF_PATCH_SIZE * std::sqrt(2.f) / 2);
    const double hp2 = HALF_PATCH_SIZE * HALF_PATCH_SIZE;
    for (v = 0; v <= vmax; ++v) o.umax[v] = ocv_round_d(std::sqrt(hp2 - v * v));
    for (v = HALF_PATCH_SIZE, v0 = 0; v >= vmin; --v) {
        while (o.umax[v0] == o.umax[v0 + 1]) ++v0;
        o.umax[v] = v0;
        ++v0;
    }
}

/* ComputePyramid, ORBextractor.cpp:1280-1304. Each level is resized from the previous level's ROI.
 * The 19-px apron is not materialised: no stage reads it (SURVEY §8a note 1). */
void compute_pyramid(Oracle& o, const uint8_t* img, int rows, int cols, size_t step) {
    o.lv.resize(o.nlevels);
    for (int l = 0; l < o.nlevels; l++) {
        Level& L = o.lv[l];
        float scale = o.isf[l];
        L.w = ocv_round_f((float)cols * scale);
        L.h = ocv_round_f((float)rows * scale);
        L.img.assign((size_t)L.w * L.h, 0);
        if (l == 0) {
            for (int y = 0; y < rows; y++) std::memcpy(&L.img[(size_t)y * cols], img + (size_t)y * step, cols);
        } else {
            const Level& P = o.lv[l - 1];
            ocv_resize_linear_u8(P.img.data(), P.w, P.h, P.w, L.img.data(), L.w, L.h, L.w);
        }
    }
}

/* ComputeKeyPointsOctTree detection part, ORBextractor.cpp:933-998: per-cell cv::FAST with threshold fallback. */
void detect_level(const Oracle& o, Level& L) {
    L.corners.clear();
    const float W = 30;
    const int minBorderX = EDGE_THRESHOLD - 3, minBorderY = minBorderX;
    const int maxBorderX = L.w - EDGE_THRESHOLD + 3, maxBorderY = L.h - EDGE_THRESHOLD + 3;
    const float width = (float)(maxBorderX - minBorderX), height = (float)(maxBorderY - minBorderY);
    const int nCols = (int)(width / W), nRows = (int)(height / W);
    if (nCols <= 0 || nRows <= 0) return;   /* reference would divide by zero; images this small are rejected upstream */
    const int wCell = (int)std::ceil(width / nCols), hCell = (int)std::ceil(height / nRows);
    std::vector<ocv_corner> buf((size_t)(wCell + 6) * (hCell + 6));
    for (int i = 0; i < nRows; i++) {
        const float iniY = (float)(minBorderY + i * hCell);
        float maxY = iniY + hCell + 6;
        if (iniY >= maxBorderY - 3) continue;
        if (maxY > maxBorderY) maxY = (float)maxBorderY;
        for (int j = 0; j < nCols; j++) {
            const float iniX = (float)(minBorderX + j * wCell);
            float maxX = iniX + wCell + 6;
            if (iniX >= maxBorderX - 6) continue;
            if (maxX > maxBorderX) maxX = (float)maxBorderX;
            const int x0 = (int)iniX, y0 = (int)iniY, cw = (int)maxX - x0, ch = (int)maxY - y0;
            const uint8_t* roi = &L.img[(size_t)y0 * L.w + x0];
            int n = ocv_fast9_16(roi, cw, ch, L.w, o.iniTh, 1, buf.data(), (int)buf.size());
            if (n == 0) n = ocv_fast9_16(roi, cw, ch, L.w, o.minTh, 1, buf.data(), (int)buf.size());
            for (int k = 0; k < n; k++) {
                /* reference: pt += (j*wCell, i*hCell) relative to (minBorderX,minBorderY); here level coords */
                Corner c = {buf[k].x + x0, buf[k].y + y0, buf[k].score};
                L.corners.push_back(c);
            }
        }
    }
}

/* ---- DistributeOctTree, ORBextractor.cpp:707-931, and ExtractorNode::DivideNode, :650-705 ----
 * The std::list is modelled as a vector of node ids in list order (front first). Keys are indices into the
 * corner list; coordinates are relative to (minX,minY) = (16,16) exactly as the reference passes them. */
struct QNode {
    int ulx, uly, brx, bry;     /* UL=(ulx,uly) UR=(brx,uly) BL=(ulx,bry) BR=(brx,bry) */
    std::vector<int> keys;      /* vKeys, original relative order preserved */
    bool noMore;
};

void divide_node(const QNode& p, const std::vector<float>& kx, const std::vector<float>& ky, QNode c[4]) {
    const int halfX = (int)std::ceil(static_cast<float>(p.brx - p.ulx) / 2);
    const int halfY = (int)std::ceil(static_cast<float>(p.bry - p.uly) / 2);
    const int mx = p.ulx + halfX, my = p.uly + halfY;
    c[0] = QNode{p.ulx, p.uly, mx, my, {}, false};
    c[1] = QNode{mx, p.uly, p.brx, my, {}, false};
    c[2] = QNode{p.ulx, my, mx, p.bry, {}, false};
    c[3] = QNode{mx, my, p.brx, p.bry, {}, false};
    for (int k : p.keys) {
        if (kx[k] < mx) { if (ky[k] < my) c[0].keys.push_back(k); else c[2].keys.push_back(k); }
        else if (ky[k] < my) c[1].keys.push_back(k);
        else c[3].keys.push_back(k);
    }
    for (int q = 0; q < 4; q++) if (c[q].keys.size() == 1) c[q].noMore = true;
}

/* returns indices (into corners) of the retained keypoints, in the reference's list order */
std::vector<int> distribute_octree(const std::vector<Corner>& corners, int minX, int maxX, int minY, int maxY, int N) {
    std::vector<float> kx(corners.size()), ky(corners.size());
    for (size_t i = 0; i < corners.size(); i++) { kx[i] = (float)(corners[i].x - minX); ky[i] = (float)(corners[i].y - minY); }

    const int nIni = (int)std::round(static_cast<float>(maxX - minX) / (maxY - minY));
    const float hX = static_cast<float>(maxX - minX) / nIni;
    std::vector<QNode> nodes;        /* storage; ids are creation order (== address order under a monotone allocator) */
    std::vector<int> order;          /* list order, front first */
    for (int i = 0; i < nIni; i++) {
        QNode n{(int)(hX * static_cast<float>(i)), 0, (int)(hX * static_cast<float>(i + 1)), maxY - minY, {}, false};
        nodes.push_back(n);
        order.push_back(i);
    }
    for (size_t i = 0; i < corners.size(); i++) nodes[(int)(kx[i] / hX)].keys.push_back((int)i);
    {
        std::vector<int> keep;
        for (int id : order) {
            if (nodes[id].keys.size() == 1) { nodes[id].noMore = true; keep.push_back(id); }
            else if (!nodes[id].keys.empty()) keep.push_back(id);
        }
        order.swap(keep);
    }

    bool finish = false;
    std::vector<std::pair<int, int>> sizeAndNode;   /* (size, node id) in creation order */
    /* split `id`: children are push_front'ed in order n1..n4 (so they end up reversed at the front) */
    auto split = [&](int id, std::vector<int>& front, int* nToExpand) {
        QNode c[4];
        divide_node(nodes[id], kx, ky, c);
        for (int q = 0; q < 4; q++) {
            if (c[q].keys.empty()) continue;
            nodes.push_back(c[q]);
            const int cid = (int)nodes.size() - 1;
            front.insert(front.begin(), cid);
            if (nodes[cid].keys.size() > 1) {
                if (nToExpand) (*nToExpand)++;
                sizeAndNode.push_back(std::make_pair((int)nodes[cid].keys.size(), cid));
            }
        }
    };

    while (!finish) {
        const int prevSize = (int)order.size();
        int nToExpand = 0;
        sizeAndNode.clear();
        std::vector<int> front, rest;
        for (int id : order) {               /* :767-830 — newly pushed nodes are not visited in this pass */
            if (nodes[id].noMore) { rest.push_back(id); continue; }
            split(id, front, &nToExpand);
        }
        order = front;
        order.insert(order.end(), rest.begin(), rest.end());

        if ((int)order.size() >= N || (int)order.size() == prevSize) {
            finish = true;
        } else if ((int)order.size() + nToExpand * 3 > N) {
            while (!finish) {                /* :841-905 */
                const int prevSize2 = (int)order.size();
                std::vector<std::pair<int, int>> prev = sizeAndNode;
                sizeAndNode.clear();
                std::sort(prev.begin(), prev.end());   /* (size, creation id) ascending == (size, address) under monotone new */
                for (int j = (int)prev.size() - 1; j >= 0; j--) {
                    const int id = prev[j].second;
                    std::vector<int> fr;
                    split(id, fr, nullptr);
                    order.erase(std::find(order.begin(), order.end(), id));
                    order.insert(order.begin(), fr.begin(), fr.end());
                    if ((int)order.size() >= N) break;
                }
                if ((int)order.size() >= N || (int)order.size() == prevSize2) finish = true;
            }
        }
    }

    std::vector<int> result;                 /* :910-928 — best response per node, first wins */
    for (int id : order) {
        const std::vector<int>& k = nodes[id].keys;
        int best = k[0];
        for (size_t i = 1; i < k.size(); i++) if (corners[k[i]].score > corners[best].score) best = k[i];
        result.push_back(best);
    }
    return result;
}

/* IC_Angle, ORBextractor.cpp:68-95 */
float ic_angle(const Level& L, float px, float py, const std::vector<int>& umax) {
    int m_01 = 0, m_10 = 0;
    const uint8_t* center = &L.img[(size_t)ocv_round_f(py) * L.w + ocv_round_f(px)];
    for (int u = -HALF_PATCH_SIZE; u <= HALF_PATCH_SIZE; ++u) m_10 += u * center[u];
    const int step = L.w;
    for (int v = 1; v <= HALF_PATCH_SIZE; ++v) {
        int v_sum = 0;
        const int d = umax[v];
        for (int u = -d; u <= d; ++u) {
            const int val_plus = center[u + v * step], val_minus = center[u - v * step];
            v_sum += (val_plus - val_minus);
            m_10 += u * (val_plus + val_minus);
        }
        m_01 += v * v_sum;
    }
    return ocv_fast_atan2((float)m_01, (float)m_10);
}

/* computeOrbDescriptor, ORBextractor.cpp:100-316: 256 steered comparisons on the blurred level */
const float factorPI = (float)(3.1415926535897932384626433832795 / 180.f);
void orb_descriptor(const Level& L, const Kp& kp, uint8_t* desc) {
    const float angle = kp.angle * factorPI;
    const float a = cosf(angle), b = sinf(angle);
    const uint8_t* center = &L.blur[(size_t)ocv_round_f(kp.y) * L.w + ocv_round_f(kp.x)];
    const int step = L.w;
    for (int i = 0; i < 32; i++) {
        int val = 0;
        for (int k = 0; k < 8; k++) {
            const Tap& t = kPattern[8 * i + k];
            const int t0 = center[ocv_round_f(t.x1 * b + t.y1 * a) * step + ocv_round_f(t.x1 * a - t.y1 * b)];
            const int t1 = center[ocv_round_f(t.x2 * b + t.y2 * a) * step + ocv_round_f(t.x2 * a - t.y2 * b)];
            val |= (t0 < t1) << k;
        }
        desc[i] = (uint8_t)val;
    }
}

/* ORBextractor::operator(), ORBextractor.cpp:1216-1278 (+ ComputeKeyPointsOctTree :999-1024) */
int extract(Oracle& o, const uint8_t* img, int rows, int cols, size_t step, Kp* out_kp, uint8_t* out_desc, int cap) {
    if (!img || rows <= 0 || cols <= 0) return 0;
    compute_pyramid(o, img, rows, cols, step);
    for (int l = 0; l < o.nlevels; l++) {
        Level& L = o.lv[l];
        detect_level(o, L);
        const int minB = EDGE_THRESHOLD - 3;
        std::vector<int> sel = distribute_octree(L.corners, minB, L.w - EDGE_THRESHOLD + 3, minB, L.h - EDGE_THRESHOLD + 3, o.quota[l]);
        const int scaledPatchSize = (int)(PATCH_SIZE * o.sf[l]);
        L.kps.clear();
        for (int idx : sel) {
            Kp k;
            k.x = (float)L.corners[idx].x; k.y = (float)L.corners[idx].y;
            k.size = (float)scaledPatchSize; k.angle = -1; k.response = (float)L.corners[idx].score; k.octave = l;
            L.kps.push_back(k);
        }
    }
    for (int l = 0; l < o.nlevels; l++)
        for (Kp& k : o.lv[l].kps) k.angle = ic_angle(o.lv[l], k.x, k.y, o.umax);
    int n = 0;
    for (int l = 0; l < o.nlevels; l++) {
        Level& L = o.lv[l];
        L.blur.clear();
        if (L.kps.empty()) continue;
        L.blur.resize(L.img.size());
        ocv_gaussian7_sigma2_u8(L.img.data(), L.w, L.h, L.w, L.blur.data(), L.w);
        const float scale = o.sf[l];
        for (const Kp& k : L.kps) {
            if (n < cap) {
                orb_descriptor(L, k, out_desc + (size_t)32 * n);
                Kp s = k;
                if (l != 0) { s.x = k.x * scale; s.y = k.y * scale; }
                out_kp[n] = s;
            }
            n++;
        }
    }
    return n;
}

}  // namespace

extern "C" {

void* orb_oracle_create(int nfeatures, float scale, int nlevels, int ini, int min) {
    Oracle* o = new Oracle();
    build_tables(*o, nfeatures, scale, nlevels, ini, min);
    return o;
}
void orb_oracle_destroy(void* p) { delete (Oracle*)p; }

/* returns the number of keypoints (may exceed cap; only cap are written). kp = ccm_keypoint[cap], desc = cap*32 */
int orb_oracle_extract(void* p, const uint8_t* img, int rows, int cols, size_t step, void* kp, uint8_t* desc, int cap) {
    return extract(*(Oracle*)p, img, rows, cols, step, (Kp*)kp, desc, cap);
}

int orb_oracle_tables(void* p, float* sf, float* isf, float* sig2, float* isig2, int* quota, int* umax16) {
    Oracle& o = *(Oracle*)p;
    for (int l = 0; l < o.nlevels; l++) { sf[l] = o.sf[l]; isf[l] = o.isf[l]; sig2[l] = o.sig2[l]; isig2[l] = o.isig2[l]; quota[l] = o.quota[l]; }
    for (int v = 0; v < 16; v++) umax16[v] = o.umax[v];
    return 0;
}

/* --- intermediate stages of the LAST extract call --- */
int orb_oracle_level_size(void* p, int level, int* w, int* h) {
    Oracle& o = *(Oracle*)p;
    if (level < 0 || level >= (int)o.lv.size()) return -1;
    *w = o.lv[level].w; *h = o.lv[level].h;
    return 0;
}
/* which: 0 = pyramid level, 1 = blurred level (empty if the level had no keypoints). Contiguous w*h. */
int orb_oracle_level_image(void* p, int level, int which, uint8_t* out, int cap) {
    Oracle& o = *(Oracle*)p;
    if (level < 0 || level >= (int)o.lv.size()) return -1;
    const std::vector<uint8_t>& v = which ? o.lv[level].blur : o.lv[level].img;
    if ((int)v.size() > cap) return -2;
    if (!v.empty()) std::memcpy(out, v.data(), v.size());
    return (int)v.size();
}
/* FAST corners handed to DistributeOctTree, reference order; xys = int[3*cap] (x, y, score), level coords */
int orb_oracle_level_corners(void* p, int level, int* xys, int cap) {
    Oracle& o = *(Oracle*)p;
    if (level < 0 || level >= (int)o.lv.size()) return -1;
    const std::vector<Corner>& c = o.lv[level].corners;
    for (int i = 0; i < (int)c.size() && i < cap; i++) { xys[3 * i] = c[i].x; xys[3 * i + 1] = c[i].y; xys[3 * i + 2] = c[i].score; }
    return (int)c.size();
}
/* keypoints of one level after distribution + orientation, list order, level (unscaled) coordinates */
int orb_oracle_level_keypoints(void* p, int level, void* kp, int cap) {
    Oracle& o = *(Oracle*)p;
    if (level < 0 || level >= (int)o.lv.size()) return -1;
    const std::vector<Kp>& k = o.lv[level].kps;
    for (int i = 0; i < (int)k.size() && i < cap; i++) ((Kp*)kp)[i] = k[i];
    return (int)k.size();
}

/* stand-alone stages, for unit parity of single kernels */
int orb_oracle_distribute(const int* xys, int n, int w, int h, int N, int* out_idx, int cap) {
    std::vector<Corner> c(n);
    for (int i = 0; i < n; i++) { c[i].x = xys[3 * i]; c[i].y = xys[3 * i + 1]; c[i].score = xys[3 * i + 2]; }
    const int minB = EDGE_THRESHOLD - 3;
    std::vector<int> sel = distribute_octree(c, minB, w - EDGE_THRESHOLD + 3, minB, h - EDGE_THRESHOLD + 3, N);
    for (int i = 0; i < (int)sel.size() && i < cap; i++) out_idx[i] = sel[i];
    return (int)sel.size();
}

}  // extern "C"
