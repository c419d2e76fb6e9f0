// Shim for the faiss header the reference's loose PQ4 copies include but do not vendor (benchmark/kernel/faiss-simd has
// no faiss tree).  TEST INFRASTRUCTURE: lets oracle/ref_wrap_pq4.cc compile the reference's pq4_fast_scan*.cpp in place.
// Only the three macros those files use; failures throw std::runtime_error like faiss's FaissException does.
#pragma once
#include <cstdio>
#include <cstring>
#include <stdexcept>
#include <string>
#define FAISS_THROW_MSG(MSG) throw std::runtime_error(std::string(MSG))
#define FAISS_THROW_IF_NOT(X) do { if (!(X)) throw std::runtime_error(std::string("Error: '" #X "' failed")); } while (0)
#define FAISS_THROW_FMT(FMT, ...) do { char b__[256]; snprintf(b__, sizeof b__, FMT, __VA_ARGS__); throw std::runtime_error(std::string(b__)); } while (0)
