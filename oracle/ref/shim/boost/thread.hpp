// stand-in for the few Boost.Thread names the CRL gui headers mention (header-level only)
#pragma once
#include <condition_variable>
#include <mutex>
#include <thread>
namespace boost {
struct mutex : std::mutex {
   typedef std::unique_lock<std::mutex> scoped_lock;
};
typedef std::condition_variable condition;
typedef std::thread thread;
}
