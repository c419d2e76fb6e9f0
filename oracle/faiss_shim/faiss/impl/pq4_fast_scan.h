// Shim (see FaissAssert.h): forwards to the reference's own copy of the header, found through -I<faiss-simd dir>.
#pragma once
#include <pq4_fast_scan.h>
