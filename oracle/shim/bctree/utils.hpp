// oracle/shim/bctree/utils.hpp -- TEST INFRASTRUCTURE ONLY: prototypes of the src/utils helpers that
// build_const_tree.cpp names (never called from interpolate()); defined as aborting stubs in ref_bctree_wrap.cpp.
#pragma once
#include <cmath>
#include <cstring>
#include <iostream>
#include <omp.h>
#include <string>
#include <sys/time.h>
#include <nlohmann/json.hpp>
std::string DateAndTime(struct timeval &tv);
void json2file(const nlohmann::json &j, const std::string &fileName);
void file2json(const std::string &fileName, nlohmann::json &j);
void *copyFile(const std::string &fileName, uint64_t size);
