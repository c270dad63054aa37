// see generated_coeffs.h in this directory
#pragma once
#include "generated_coeffs.h"
