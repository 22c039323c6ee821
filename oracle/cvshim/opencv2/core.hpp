/* oracle/cvshim — TEST INFRASTRUCTURE.  Stand-in for the OpenCV header; everything lives in opencv2/opencv.hpp. */
#pragma once
#include "opencv2/opencv.hpp"
