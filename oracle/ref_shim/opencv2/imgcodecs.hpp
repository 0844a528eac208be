// TEST INFRASTRUCTURE ONLY -- see shim.hpp
#pragma once
#include <opencv2/shim.hpp>
