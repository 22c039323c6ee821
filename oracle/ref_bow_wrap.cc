/*
 * oracle/ref_bow_wrap.cc — TEST INFRASTRUCTURE ONLY.
 *
 * C entry points around the reference's VENDORED, UNMODIFIED DBoW2 (/root/reference/Thirdparty/DBoW2/DBoW2/*.cpp,
 * TemplatedVocabulary.h, DUtils/Random.cpp, Timestamp.cpp), compiled where it lies by oracle/build_ref.sh against oracle/cvshim
 * and oracle/boostshim into oracle/_ref/libref_bow.so.  Pins orc_bow_transform / orc_bow_vectors (oracle/orc_match.c) and the
 * text-file reader of orb-slam3_b200/cpp/ORBVocabulary.h to the code ORB-SLAM3 runs in Frame::ComputeBoW (src/Frame.cc:905-922):
 *   ORBVocabulary = TemplatedVocabulary<FORB::TDescriptor, FORB>          include/ORBVocabulary.h:29-30
 *   loadFromTextFile                                                      TemplatedVocabulary.h:1338-1415
 *   transform(features, BowVector&, FeatureVector&, levelsup = 4)         TemplatedVocabulary.h:1174-1216
 *   transform(feature, WordId&, WordValue&, NodeId*, levelsup)            TemplatedVocabulary.h:1218-1258 (protected)
 */
#define protected public
#include "DBoW2/FORB.h"
#include "DBoW2/TemplatedVocabulary.h"
#undef protected

#include <cstdint>
#include <cstring>

typedef DBoW2::TemplatedVocabulary<DBoW2::FORB::TDescriptor, DBoW2::FORB> RefVocabulary;

extern "C" {

void* ref_voc_load_text(const char* path)
{
    RefVocabulary* v = new RefVocabulary();
    if (!v->loadFromTextFile(path)) { delete v; return nullptr; }
    return v;
}

void ref_voc_destroy(void* v) { delete static_cast<RefVocabulary*>(v); }

int ref_voc_info(void* vp, int* k, int* L, int* nnodes, int* nwords)
{
    RefVocabulary* v = static_cast<RefVocabulary*>(vp);
    *k = v->getBranchingFactor(); *L = v->getDepthLevels(); *nwords = (int)v->size(); *nnodes = (int)v->m_nodes.size();
    return 0;
}

/* per feature: (word id, weight, node id at L - levelsup) */
void ref_voc_transform_each(void* vp, const uint8_t* desc, int n, int levelsup, int32_t* word, double* weight, int32_t* node)
{
    RefVocabulary* v = static_cast<RefVocabulary*>(vp);
    for (int i = 0; i < n; i++) {
        cv::Mat d(1, 32, CV_8U);
        std::memcpy(d.data, desc + 32 * (size_t)i, 32);
        DBoW2::WordId id; DBoW2::WordValue w; DBoW2::NodeId nid = 0;   /* left untouched when the word is shallower than L - levelsup */
        v->transform(d, id, w, &nid, levelsup);
        word[i] = (int32_t)id; weight[i] = (double)w; node[i] = (int32_t)nid;
    }
}

/* Frame::ComputeBoW: the BowVector (ascending word id) and the FeatureVector (ascending node id, feature indices in push order) */
int ref_voc_compute_bow(void* vp, const uint8_t* desc, int n, int levelsup, int32_t* bow_id, double* bow_val,
                        int32_t* fv_node, int32_t* fv_start, int32_t* fv_idx, int32_t* nfv)
{
    RefVocabulary* v = static_cast<RefVocabulary*>(vp);
    std::vector<cv::Mat> feats; feats.reserve((size_t)n);
    for (int i = 0; i < n; i++) {                       /* Converter::toDescriptorVector, src/Converter.cc:28-36: one row each */
        cv::Mat d(1, 32, CV_8U);
        std::memcpy(d.data, desc + 32 * (size_t)i, 32);
        feats.push_back(d);
    }
    DBoW2::BowVector bv; DBoW2::FeatureVector fv;
    v->transform(feats, bv, fv, levelsup);
    int nb = 0;
    for (DBoW2::BowVector::const_iterator it = bv.begin(); it != bv.end(); ++it, ++nb) { bow_id[nb] = (int32_t)it->first; bow_val[nb] = it->second; }
    int k = 0, p = 0;
    for (DBoW2::FeatureVector::const_iterator it = fv.begin(); it != fv.end(); ++it, ++k) {
        fv_node[k] = (int32_t)it->first; fv_start[k] = p;
        for (size_t j = 0; j < it->second.size(); j++) fv_idx[p++] = (int32_t)it->second[j];
    }
    fv_start[k] = p; *nfv = k;
    return nb;
}

/* BowVector scoring used by the place-recognition database (L1, ScoringObject.cpp:32-63) */
double ref_voc_score(void* vp, const int32_t* id1, const double* val1, int n1, const int32_t* id2, const double* val2, int n2)
{
    RefVocabulary* v = static_cast<RefVocabulary*>(vp);
    DBoW2::BowVector a, b;
    for (int i = 0; i < n1; i++) a.addWeight((DBoW2::WordId)id1[i], val1[i]);
    for (int i = 0; i < n2; i++) b.addWeight((DBoW2::WordId)id2[i], val2[i]);
    return v->score(a, b);
}

} // extern "C"
