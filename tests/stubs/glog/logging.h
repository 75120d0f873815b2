// Test stub: LOG(severity) << ... to stderr.
#pragma once
#include <iostream>
#include <sstream>
namespace stub_glog {
struct Line {
    explicit Line(const char* sev) { s << "[" << sev << "] "; }
    ~Line() { std::cerr << s.str() << std::endl; }
    std::ostringstream s;
};
}  // namespace stub_glog
#define LOG(sev) ::stub_glog::Line(#sev).s
