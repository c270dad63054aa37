// scheme/ckksrns/ckksrns-fhe.h -- forwarding header (see openfhe.h in this directory)
#pragma once
#include "../../openfhe.h"
