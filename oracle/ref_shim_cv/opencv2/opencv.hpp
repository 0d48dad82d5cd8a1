#pragma once
#include "../opencv/cv.h"
