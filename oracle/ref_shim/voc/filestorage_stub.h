// TEST INFRASTRUCTURE (oracle/_ref/libref_voc.so). The vendored DBoW2 TemplatedVocabulary.h names cv::FileStorage /
// cv::FileNode in its YAML save()/load() members (virtual, so they are instantiated with the class). The parity
// harness loads vocabularies through the fork's loadFromTextFile (TemplatedVocabulary.h:1362-1436) only; these types
// exist so that the header compiles and abort if the YAML path is ever reached.
#pragma once
#include <cstdlib>
#include <sstream>
#include <string>

#include <opencv2/core/core.hpp>

namespace cv {
struct FileNode {
  FileNode operator[](const char*) const { std::abort(); }
  FileNode operator[](const std::string&) const { std::abort(); }
  FileNode operator[](int) const { std::abort(); }
  FileNode operator[](unsigned) const { std::abort(); }
  size_t size() const { std::abort(); }
  operator int() const { std::abort(); }
  operator double() const { std::abort(); }
  operator std::string() const { std::abort(); }
};
struct FileStorage {
  enum { READ = 0, WRITE = 1 };
  FileStorage(const char*, int) {}
  bool isOpened() const { return false; }
  FileNode operator[](const std::string&) const { std::abort(); }
  void release() {}
};
template <class T>
inline FileStorage& operator<<(FileStorage& fs, const T&) { std::abort(); return fs; }
}  // namespace cv
