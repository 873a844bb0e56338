// minicv shim: forwards to the single shim header (oracle/minicv/minicv.hpp).
#pragma once
// the real opencv2/core/core.hpp pulls these in transitively, and DBoW2's TemplatedVocabulary.h relies on it
// (std::stringstream / std::cerr without their headers, Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1350)
#include <iostream>
#include <sstream>
#include "minicv.hpp"
