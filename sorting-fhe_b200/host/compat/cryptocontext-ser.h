// cryptocontext-ser.h -- forwarding header: Serial / SerType over the engine (host/serial.h)
#pragma once
#include "../serial.h"
