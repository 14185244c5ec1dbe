// Stand-in for <rclcpp/rclcpp.hpp> (ROS 2 Humble is not in this image).  The reference's frontier search only asks whether the
// process is still up (DEP/src/FrontierSearch.cpp:58,123); FisherInfoManager.cpp creates a node, a service client for
// slam_msgs/srv/GetLandmarksInView and a publisher, and waits for the service's answer: here a client answers at once through a
// handler the harness installs (oracle/ref_harness.cpp).  Test infrastructure for oracle/_ref, never shipped.
#pragma once
// (the real header brings these standard headers in; the reference's sources rely on that)
#include <algorithm>
#include <array>
#include <chrono>
#include <functional>
#include <limits>
#include <memory>
#include <mutex>
#include <queue>
#include <string>
#include <vector>
namespace rclcpp {
inline bool ok() { return true; }
inline void shutdown() {}
enum class FutureReturnCode { SUCCESS, INTERRUPTED, TIMEOUT };

template <typename ServiceT>
class Client {
public:
  using SharedPtr = std::shared_ptr<Client<ServiceT>>;
  using Handler = std::function<void(const typename ServiceT::Request &, typename ServiceT::Response &)>;
  struct SharedFuture {
    std::shared_ptr<typename ServiceT::Response> response;
    std::shared_ptr<typename ServiceT::Response> get() const { return response; }
  };
  static Handler &handler() { static Handler h; return h; }
  template <typename Duration> bool wait_for_service(Duration) { return static_cast<bool>(handler()); }
  SharedFuture async_send_request(std::shared_ptr<typename ServiceT::Request> request) {
    SharedFuture f;
    f.response = std::make_shared<typename ServiceT::Response>();
    handler()(*request, *f.response);
    return f;
  }
};
template <typename MessageT>
class Publisher {
public:
  using SharedPtr = std::shared_ptr<Publisher<MessageT>>;
  void publish(const MessageT &) {}
};
class Node {
public:
  using SharedPtr = std::shared_ptr<Node>;
  explicit Node(const std::string & = "") {}
  static SharedPtr make_shared(const std::string &name) { return std::make_shared<Node>(name); }
  template <typename ServiceT> typename Client<ServiceT>::SharedPtr create_client(const std::string &) { return std::make_shared<Client<ServiceT>>(); }
  template <typename MessageT> typename Publisher<MessageT>::SharedPtr create_publisher(const std::string &, int) { return std::make_shared<Publisher<MessageT>>(); }
};
namespace executors { class MultiThreadedExecutor {}; }
template <typename FutureT>
FutureReturnCode spin_until_future_complete(Node::SharedPtr, const FutureT &) { return FutureReturnCode::SUCCESS; }
}  // namespace rclcpp
