/* see application.h (shim) */
#pragma once
#include "application.h"
