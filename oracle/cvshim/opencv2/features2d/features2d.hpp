/* oracle/cvshim — TEST INFRASTRUCTURE.  Stand-in for the OpenCV header the unmodified reference
 * sources include; everything lives in opencv2/opencv.hpp. */
#pragma once
#include "opencv2/opencv.hpp"
