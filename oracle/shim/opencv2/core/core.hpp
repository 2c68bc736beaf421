/*
 * Stand-in for <opencv2/core/core.hpp> -- TEST INFRASTRUCTURE, used only by oracle/Makefile to compile the reference's
 * own DBoW2 sources where they lie (/root/reference/Thirdparty/DBoW2) into oracle/_ref/libdbow2_ref.so.  This image has
 * no OpenCV C++ headers; the vocabulary path (loadFromTextFile / create / transform / score) touches cv::Mat only as a
 * 1x32 byte container (create, ptr<>, clone, zeros, cols), which include/orbb200/cv_shim.h already provides.  The
 * cv::FileStorage members (yml save/load) are declared so that the template parses; they are never called and a call
 * would fail at link time, not silently.
 */
#ifndef ORBB200_ORACLE_OPENCV_CORE_STANDIN
#define ORBB200_ORACLE_OPENCV_CORE_STANDIN
#include <math.h>
#include <stdlib.h>
#include <algorithm>
#include <cmath>
#include <iostream>
#include <sstream> /* the real header drags these in and the DBoW2 sources rely on it */
#include <string>

#include "orbb200/cv_shim.h"

namespace cv {
class FileNode {
public:
    size_t size() const;
    FileNode operator[](const char *) const;
    FileNode operator[](const std::string &) const;
    FileNode operator[](int) const;
    operator int() const;
    operator double() const;
    operator std::string() const;
};
class FileStorage {
public:
    enum { READ = 0, WRITE = 1 };
    FileStorage(const char *, int);
    FileStorage(const std::string &, int);
    bool isOpened() const;
    void release();
    FileNode operator[](const std::string &) const;
    FileNode operator[](const char *) const;
};
template <typename T> FileStorage &operator<<(FileStorage &fs, const T &);
} // namespace cv
#endif
