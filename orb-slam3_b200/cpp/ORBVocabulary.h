// ORBVocabulary.h — device-resident stand-in for the reference's include/ORBVocabulary.h
// (typedef DBoW2::TemplatedVocabulary<DBoW2::FORB::TDescriptor, DBoW2::FORB> ORBVocabulary) on the C ABI:
// loadFromTextFile parses ORBvoc.txt exactly as Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1338-1415 does and
// uploads the tree; transform() is Frame::ComputeBoW's call (src/Frame.cc:905-922) with the BowVector /
// FeatureVector returned flat (ascending word ids with L1-normalised values; CSR of node -> feature indices), which
// is also the layout ORBmatcherGPU::KeyFrameView::SetFeatVec / orbx_search_by_bow take.
#pragma once
#include <cstdint>
#include <fstream>
#include <sstream>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/orbx.h"

namespace ORB_SLAM3 {

class ORBVocabulary {
public:
    explicit ORBVocabulary(int device = 0) : device_(device) {}
    ~ORBVocabulary() { orbx_vocabulary_destroy(h_); }
    ORBVocabulary(const ORBVocabulary&) = delete;
    ORBVocabulary& operator=(const ORBVocabulary&) = delete;

    // bool TemplatedVocabulary::loadFromTextFile(const std::string&), TemplatedVocabulary.h:1338-1415
    bool loadFromTextFile(const std::string& filename)
    {
        std::ifstream f(filename.c_str());
        if (!f.is_open() || f.eof()) return false;
        std::string s;
        std::getline(f, s);
        std::stringstream ss(s);
        int n1 = 0, n2 = 0;
        ss >> k_ >> L_ >> n1 >> n2;
        if (k_ < 0 || k_ > 20 || L_ < 1 || L_ > 10 || n1 < 0 || n1 > 5 || n2 < 0 || n2 > 3) return false;    // :1359
        if (n1 != 0 || n2 != 0) throw std::runtime_error("ORBVocabulary: only L1_NORM scoring / TF_IDF weighting (ORBvoc.txt) is implemented");
        std::vector<int32_t> parent; std::vector<uint8_t> leaf, desc; std::vector<double> weight;
        while (std::getline(f, s)) {
            if (s.empty()) continue;      // the reference's `while(!f.eof())` turns a trailing empty line into one more node built from
                                          // uninitialised locals (pid, nIsLeaf of the previous row in practice); not reproduced
            std::stringstream sn(s);
            int pid = 0, isLeaf = 0; double w = 0;
            sn >> pid >> isLeaf;
            for (int i = 0; i < 32; i++) { int b = 0; sn >> b; desc.push_back((uint8_t)b); }               // FORB::fromString
            sn >> w;
            parent.push_back(pid); leaf.push_back(isLeaf > 0); weight.push_back(w);
        }
        if (parent.empty()) return false;
        orbx_vocabulary_destroy(h_); h_ = nullptr;
        if (orbx_vocabulary_create(device_, k_, L_, (int)parent.size(), parent.data(), leaf.data(), desc.data(), weight.data(), &h_) != ORBX_OK)
            throw std::runtime_error(std::string("ORBVocabulary: ") + orbx_last_error());
        return true;
    }

    struct BowVectorFlat { std::vector<int32_t> id; std::vector<double> value; };                 // DBoW2::BowVector in map order
    struct FeatureVectorFlat { std::vector<int32_t> node, start, idx; };                          // DBoW2::FeatureVector as CSR

    // void transform(const std::vector<TDescriptor>& features, BowVector& v, FeatureVector& fv, int levelsup) :1127-1196
    // descriptors: n rows of 32 bytes (mDescriptors.data when continuous).
    void transform(const uint8_t* descriptors, int n, BowVectorFlat& v, FeatureVectorFlat& fv, int levelsup = 4) const
    {
        v.id.assign((size_t)n, 0); v.value.assign((size_t)n, 0.0);
        fv.node.assign((size_t)n, 0); fv.start.assign((size_t)n + 1, 0); fv.idx.assign((size_t)n, 0);
        int nb = 0, nf = 0;
        if (orbx_compute_bow(h_, descriptors, n, levelsup, v.id.data(), v.value.data(), &nb, fv.node.data(), fv.start.data(), fv.idx.data(), &nf) != ORBX_OK)
            throw std::runtime_error(std::string("ORBVocabulary::transform: ") + orbx_last_error());
        v.id.resize((size_t)nb); v.value.resize((size_t)nb);
        fv.node.resize((size_t)nf); fv.start.resize((size_t)nf + 1); fv.idx.resize((size_t)fv.start[(size_t)nf]);
    }

    bool empty() const { return h_ == nullptr; }
    const orbx_vocabulary* handle() const { return h_; }

private:
    int device_, k_ = 0, L_ = 0;
    orbx_vocabulary* h_ = nullptr;
};

}  // namespace ORB_SLAM3
