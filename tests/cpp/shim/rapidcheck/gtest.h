// <rapidcheck/gtest.h> stand-in: RC_GTEST_PROP lives in ../rapidcheck.h
#pragma once
#include "../rapidcheck.h"
