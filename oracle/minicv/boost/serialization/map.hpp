// minicv shim (TEST INFRASTRUCTURE): see serialization.hpp
#pragma once
#include "serialization.hpp"
