// minicv shim: forwards to the single shim header (oracle/minicv/minicv.hpp).
#pragma once
#include "minicv.hpp"
