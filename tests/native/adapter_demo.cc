// Compile / run check of the C++ drop-in adapters against the cv:: shim: mirrors the reference's own call sites
// (Tracking: new ORBextractor(...), Frame::ExtractORB: (*extractor)(im, cv::Mat(), keys, desc, vLapping)).
// argv[1] = raw 8-bit image file, argv[2] = width, argv[3] = height.  Prints "mono n checksum".
#include "cv_shim.h"
#include "../../mam3slam_b200/cpp/ORBextractor.h"
#include "../../mam3slam_b200/cpp/ORBmatcher.h"
#include <cstdio>
#include <cstdlib>

int main(int argc, char** argv) {
    if (argc < 4) return 2;
    const int w = atoi(argv[2]), h = atoi(argv[3]);
    std::vector<uint8_t> px((size_t)w * h);
    FILE* f = fopen(argv[1], "rb");
    if (!f || fread(px.data(), 1, px.size(), f) != px.size()) return 3;
    fclose(f);
    cv::Mat im(h, w, CV_8UC1, px.data(), (size_t)w);
    MAM3SLAM::ORBextractor* mpORBextractorLeft = new MAM3SLAM::ORBextractor(1000, 1.2f, 8, 20, 7);
    std::vector<cv::KeyPoint> mvKeys;
    cv::Mat mDescriptors;
    std::vector<int> vLapping = {0, 1000};
    const int monoLeft = (*mpORBextractorLeft)(im, cv::Mat(), mvKeys, mDescriptors, vLapping);
    unsigned long long cs = 0;
    for (size_t i = 0; i < mvKeys.size(); ++i) {
        cs = cs * 1315423911ull + (unsigned)mvKeys[i].pt.x * 31u + (unsigned)mvKeys[i].pt.y * 7u + (unsigned)mvKeys[i].octave;
        for (int b = 0; b < 32; ++b) cs = cs * 131ull + mDescriptors.ptr((int)i)[b];
    }
    int d01 = mvKeys.size() > 1 ? MAM3SLAM::ORBmatcher::DescriptorDistance(mDescriptors.row(0), mDescriptors.row(1)) : -1;
    mpORBextractorLeft->DownloadImagePyramid();
    printf("%d %zu %llu %d %d\n", monoLeft, mvKeys.size(), cs, d01, mpORBextractorLeft->mvImagePyramid[7].cols);
    delete mpORBextractorLeft;
    return 0;
}
