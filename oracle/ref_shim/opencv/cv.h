// ORACLE / test infrastructure.  Minimal stand-in for <opencv/cv.h>, just enough for the reference's DBoW2 sources
// (Thirdparty/DBoW2/DBoW2/{TemplatedVocabulary.h,FORB.cpp,...}) to compile where they lie, WITHOUT OpenCV:
//   cv::Mat          - a reference-counted 2-D byte/float matrix (create / ptr / clone / zeros / release), which is all
//                      FORB.cpp asks of it;
//   cv::FileStorage / cv::FileNode - inert: the YAML save/load members of TemplatedVocabulary are virtual and so get
//                      instantiated, but oracle/_ref only ever uses loadFromTextFile / saveToTextFile, which do not
//                      touch them.  isOpened() is false, so the YAML entry points throw instead of silently doing
//                      nothing.
// Nothing here restates OpenCV behaviour that the BoW path's results depend on (FORB::distance reads raw bytes).
#pragma once
// the real header drags these in; the DBoW2 sources rely on that (unqualified pow/log, stringstream)
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <algorithm>
#include <iostream>
#include <limits>
#include <sstream>
#include <cstdint>
#include <cstring>
#include <memory>
#include <string>
#include <vector>

#define CV_8U 0
#define CV_32F 5

namespace cv {

class Mat {
 public:
  int rows = 0, cols = 0;
  Mat() = default;
  Mat(int r, int c, int type) { create(r, c, type); }
  void create(int r, int c, int type) {
    const size_t es = type == CV_32F ? 4 : 1;
    if (buf_ && r == rows && c == cols && es == esz_) return;
    rows = r; cols = c; esz_ = es;
    buf_ = std::make_shared<std::vector<unsigned char>>((size_t)r * c * es);
  }
  void release() { buf_.reset(); rows = cols = 0; }
  bool empty() const { return !buf_ || rows * cols == 0; }
  Mat clone() const {
    Mat m;
    m.rows = rows; m.cols = cols; m.esz_ = esz_;
    if (buf_) m.buf_ = std::make_shared<std::vector<unsigned char>>(*buf_);
    return m;
  }
  static Mat zeros(int r, int c, int type) { return Mat(r, c, type); }  // vector storage is value-initialised
  template <class T> T* ptr(int row = 0) { return reinterpret_cast<T*>(buf_->data() + (size_t)row * cols * esz_); }
  template <class T> const T* ptr(int row = 0) const {
    return reinterpret_cast<const T*>(buf_->data() + (size_t)row * cols * esz_);
  }
  Mat row(int r) const {  // deep copy of one row (enough for descriptor vectors)
    Mat m;
    m.create(1, cols, esz_ == 4 ? CV_32F : CV_8U);
    std::memcpy(m.buf_->data(), buf_->data() + (size_t)r * cols * esz_, (size_t)cols * esz_);
    return m;
  }

 private:
  size_t esz_ = 1;
  std::shared_ptr<std::vector<unsigned char>> buf_;
};

class FileNode {
 public:
  size_t size() const { return 0; }
  FileNode operator[](int) const { return FileNode(); }
  FileNode operator[](const char*) const { return FileNode(); }
  FileNode operator[](const std::string&) const { return FileNode(); }
  operator int() const { return 0; }
  operator float() const { return 0.f; }
  operator double() const { return 0.0; }
  operator std::string() const { return std::string(); }
};

class FileStorage {
 public:
  enum { READ = 0, WRITE = 1 };
  FileStorage(const std::string&, int) {}
  bool isOpened() const { return false; }
  FileNode operator[](const char*) const { return FileNode(); }
  FileNode operator[](const std::string&) const { return FileNode(); }
};
template <class T> inline FileStorage& operator<<(FileStorage& fs, const T&) { return fs; }

}  // namespace cv
