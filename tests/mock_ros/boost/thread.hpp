#pragma once
#include <thread>
namespace boost { using thread = std::thread; }
