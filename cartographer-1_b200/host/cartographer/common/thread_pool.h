// common::Task / common::ThreadPoolInterface / common::ThreadPool with the
// reference's public interface (cartographer/common/task.h:30-80,
// cartographer/common/thread_pool.h:35-90): tasks with dependencies, scheduled on a
// fixed set of worker threads; a task runs once all its dependencies completed.
// The constraint builder only needs Schedule(), SetWorkItem(), AddDependency() and
// GetState(); in a full Cartographer tree the real headers replace this file.
#ifndef CARTOGRAPHER_COMMON_THREAD_POOL_H_
#define CARTOGRAPHER_COMMON_THREAD_POOL_H_

#include <condition_variable>
#include <deque>
#include <functional>
#include <map>
#include <memory>
#include <mutex>
#include <set>
#include <thread>
#include <vector>

namespace cartographer {
namespace common {

class ThreadPoolInterface;

class Task {
 public:
  friend class ThreadPoolInterface;
  using WorkItem = std::function<void()>;
  enum State { NEW, DISPATCHED, DEPENDENCIES_COMPLETED, RUNNING, COMPLETED };

  Task() = default;
  ~Task();

  State GetState();
  // State must be 'NEW'.
  void SetWorkItem(const WorkItem& work_item);
  // State must be 'NEW'.  An expired 'dependency' counts as completed.
  void AddDependency(std::weak_ptr<Task> dependency);

 private:
  void AddDependentTask(Task* dependent_task);
  void Execute();
  void SetThreadPool(ThreadPoolInterface* thread_pool);
  void OnDependenyCompleted();

  WorkItem work_item_;
  ThreadPoolInterface* thread_pool_to_notify_ = nullptr;
  State state_ = NEW;
  unsigned int uncompleted_dependencies_ = 0;
  std::set<Task*> dependent_tasks_;
  std::mutex mutex_;
};

class ThreadPoolInterface {
 public:
  ThreadPoolInterface() {}
  virtual ~ThreadPoolInterface() {}
  virtual std::weak_ptr<Task> Schedule(std::unique_ptr<Task> task) = 0;

 protected:
  void Execute(Task* task);
  void SetThreadPool(Task* task);

 private:
  friend class Task;
  virtual void NotifyDependenciesCompleted(Task* task) = 0;
};

class ThreadPool : public ThreadPoolInterface {
 public:
  explicit ThreadPool(int num_threads);
  ~ThreadPool();
  ThreadPool(const ThreadPool&) = delete;
  ThreadPool& operator=(const ThreadPool&) = delete;

  std::weak_ptr<Task> Schedule(std::unique_ptr<Task> task) override;

 private:
  void DoWork();
  void NotifyDependenciesCompleted(Task* task) override;

  std::mutex mutex_;
  std::condition_variable wake_;
  bool running_ = true;
  std::vector<std::thread> pool_;
  std::deque<std::shared_ptr<Task>> task_queue_;
  std::map<Task*, std::shared_ptr<Task>> tasks_not_ready_;
};

}  // namespace common
}  // namespace cartographer

#endif  // CARTOGRAPHER_COMMON_THREAD_POOL_H_
