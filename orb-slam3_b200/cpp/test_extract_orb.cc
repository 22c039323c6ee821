// test_extract_orb.cc — the ORBX_WITH_OPENCV branch of cpp/ORBextractor.h compiled against an OpenCV-shaped cv:: namespace
// (tests build it with -DORBX_WITH_OPENCV -Ioracle/cvshim; with a real OpenCV the same source builds against it), driven by
// the caller the reference has: Frame::ExtractORB (src/Frame.cc:482-492) on extractor objects built as in
// src/Tracking.cc:1106-1113.  Writes what the frame ends up holding to a file the test compares with the oracle.
//   test_extract_orb <frame.raw> <rows> <cols> <row step> <nfeatures> <x0> <x1> <out.bin>
#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <iterator>

#include "ORBextractor.h"

using namespace std;

namespace ORB_SLAM3 {
// the members of Frame that ExtractORB touches (include/Frame.h:160-189, 226-228)
struct Frame {
    ORBextractor *mpORBextractorLeft, *mpORBextractorRight;
    std::vector<cv::KeyPoint> mvKeys, mvKeysRight;
    cv::Mat mDescriptors, mDescriptorsRight;
    int monoLeft, monoRight;
    // src/Frame.cc:482-492
    void ExtractORB(int flag, const cv::Mat& im, const int x0, const int x1)
    {
        vector<int> vLapping = {x0, x1};
        if (flag == 0)
            monoLeft = (*mpORBextractorLeft)(im, cv::Mat(), mvKeys, mDescriptors, vLapping);
        else
            monoRight = (*mpORBextractorRight)(im, cv::Mat(), mvKeysRight, mDescriptorsRight, vLapping);
    }
};
}  // namespace ORB_SLAM3

int main(int argc, char** argv)
{
    if (argc != 9) { fprintf(stderr, "usage: test_extract_orb frame.raw rows cols step nfeatures x0 x1 out.bin\n"); return 2; }
    const int rows = atoi(argv[2]), cols = atoi(argv[3]), step = atoi(argv[4]), nfeatures = atoi(argv[5]), x0 = atoi(argv[6]), x1 = atoi(argv[7]);
    ifstream in(argv[1], ios::binary);
    vector<unsigned char> buf((istreambuf_iterator<char>(in)), istreambuf_iterator<char>());
    if ((int)buf.size() < rows * step) { fprintf(stderr, "short frame file\n"); return 2; }
    try {
        cv::Mat im(rows, cols, CV_8UC1, buf.data(), (size_t)step);           // a view with its own row step, as a ROI of a larger image would be
        ORB_SLAM3::Frame F;
        // src/Tracking.cc:1106-1108: new ORBextractor(nFeatures, fScaleFactor, nLevels, fIniThFAST, fMinThFAST)
        F.mpORBextractorLeft = new ORB_SLAM3::ORBextractor(nfeatures, 1.2f, 8, 20, 7);
        F.mpORBextractorRight = new ORB_SLAM3::ORBextractor(nfeatures, 1.2f, 8, 20, 7);
        F.mpORBextractorRight->SetKeepHostPyramid(true);
        F.ExtractORB(0, im, x0, x1);
        F.ExtractORB(1, im, x0, x1);
        // the InputArray / OutputArray form of the reference's signature (include/ORBextractor.h:57-59)
        vector<cv::KeyPoint> k3; cv::Mat d3; vector<int> lap = {x0, x1};
        cv::InputArray ia = im; cv::OutputArray oa = d3;
        const int mono3 = (*F.mpORBextractorLeft)(ia, cv::noArray(), k3, oa, lap);
        bool same = mono3 == F.monoLeft && F.monoRight == F.monoLeft && k3.size() == F.mvKeys.size() && F.mvKeysRight.size() == F.mvKeys.size() &&
                    d3.rows == F.mDescriptors.rows && d3.cols == 32 && d3.type() == CV_8U;
        for (size_t i = 0; same && i < k3.size(); i++)
            same = memcmp(&k3[i], &F.mvKeys[i], sizeof(cv::KeyPoint)) == 0 && memcmp(&F.mvKeysRight[i], &F.mvKeys[i], sizeof(cv::KeyPoint)) == 0 &&
                   memcmp(d3.ptr((int)i), F.mDescriptors.ptr((int)i), 32) == 0 && memcmp(F.mDescriptorsRight.ptr((int)i), F.mDescriptors.ptr((int)i), 32) == 0;
        if (!same) { fprintf(stderr, "the three call forms disagree\n"); return 3; }
        // an empty image returns -1 and leaves the outputs alone (:1538)
        vector<cv::KeyPoint> k4; cv::Mat d4;
        if ((*F.mpORBextractorLeft)(cv::Mat(), cv::Mat(), k4, d4, lap) != -1) { fprintf(stderr, "empty image did not return -1\n"); return 3; }
        ofstream out(argv[8], ios::binary);
        const int n = (int)F.mvKeys.size(), mono = F.monoLeft, nlev = (int)F.mpORBextractorRight->mvImagePyramid.size();
        out.write((const char*)&mono, 4); out.write((const char*)&n, 4); out.write((const char*)&nlev, 4);
        out.write((const char*)F.mvKeys.data(), (streamsize)(sizeof(cv::KeyPoint) * (size_t)n));
        for (int i = 0; i < n; i++) out.write((const char*)F.mDescriptors.ptr(i), 32);
        for (int l = 0; l < nlev; l++) {                                     // mvImagePyramid kept on the host on request
            const cv::Mat& m = F.mpORBextractorRight->mvImagePyramid[(size_t)l];
            out.write((const char*)&m.rows, 4); out.write((const char*)&m.cols, 4);
            for (int y = 0; y < m.rows; y++) out.write((const char*)m.ptr(y), m.cols);
        }
        printf("ok mono %d n %d levels %d\n", mono, n, nlev);
        return 0;
    } catch (const std::exception& e) {
        fprintf(stderr, "error: %s\n", e.what());
        return 1;
    }
}
