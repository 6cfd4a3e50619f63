/* TEST INFRASTRUCTURE ONLY -- force-included (-include) ahead of the reference sources when building
 * oracle/_ref/libnsref.so: pull in every standard header first (their include guards keep them intact),
 * then make the reference classes' private members reachable from ref_driver.cpp. */
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <iostream>
#include <iomanip>
#include <fstream>
#include <sstream>
#include <string>
#include <vector>
#include <set>
#include <algorithm>
#include <chrono>
#include <utility>
#define private public
