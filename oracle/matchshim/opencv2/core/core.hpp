/* TEST INFRASTRUCTURE — <opencv2/core/core.hpp> as the vendored DBoW2 headers include it: the mini cv::Mat of
 * opencv2/opencv.hpp plus inert cv::FileStorage / cv::FileNode declarations, enough for the (virtual, hence always
 * instantiated) YAML save/load members of TemplatedVocabulary.h to compile. They are never called: the vocabulary is
 * loaded through loadFromTextFile, which uses iostreams only. */
#ifndef MATCHSHIM_OPENCV_CORE_HPP_
#define MATCHSHIM_OPENCV_CORE_HPP_
#include <iostream>
#include <sstream>   /* the real header pulls these in; TemplatedVocabulary.h relies on it */
#include <stdexcept>
#include <string>
#include <opencv2/opencv.hpp>

namespace cv {
struct FileNode {
    FileNode operator[](const char*) const { throw std::logic_error("matchshim: cv::FileNode is inert"); }
    FileNode operator[](const std::string&) const { throw std::logic_error("matchshim: cv::FileNode is inert"); }
    FileNode operator[](int) const { throw std::logic_error("matchshim: cv::FileNode is inert"); }
    size_t size() const { return 0; }
    operator int() const { return 0; }
    operator double() const { return 0; }
    operator std::string() const { return std::string(); }
};
struct FileStorage {
    enum { READ = 0, WRITE = 1 };
    FileStorage(const char*, int) {}
    bool isOpened() const { return false; }
    FileNode operator[](const char*) const { return FileNode(); }
    FileNode operator[](const std::string&) const { return FileNode(); }
};
template <typename T> inline FileStorage& operator<<(FileStorage& fs, const T&) { return fs; }
}  // namespace cv
#endif
