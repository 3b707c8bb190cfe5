#include "cartographer/common/thread_pool.h"

#include <unistd.h>

#include "cartographer_compat.h"

namespace cartographer {
namespace common {

// ---- Task (reference semantics: cartographer/common/task.cc) -------------------
Task::~Task() {
  // A task must not be destroyed while dispatched but unfinished.
  if (state_ != NEW && state_ != COMPLETED) {
    std::fprintf(stderr, "common::Task destroyed in state %d\n", static_cast<int>(state_));
  }
}

Task::State Task::GetState() {
  std::lock_guard<std::mutex> lock(mutex_);
  return state_;
}

void Task::SetWorkItem(const WorkItem& work_item) {
  std::lock_guard<std::mutex> lock(mutex_);
  CSM_CHECK(state_ == NEW);
  work_item_ = work_item;
}

void Task::AddDependency(std::weak_ptr<Task> dependency) {
  std::shared_ptr<Task> shared;
  {
    std::lock_guard<std::mutex> lock(mutex_);
    CSM_CHECK(state_ == NEW);
    if ((shared = dependency.lock())) ++uncompleted_dependencies_;
  }
  if (shared) shared->AddDependentTask(this);
}

void Task::SetThreadPool(ThreadPoolInterface* thread_pool) {
  std::lock_guard<std::mutex> lock(mutex_);
  CSM_CHECK(state_ == NEW);
  state_ = DISPATCHED;
  thread_pool_to_notify_ = thread_pool;
  if (uncompleted_dependencies_ == 0) {
    state_ = DEPENDENCIES_COMPLETED;
    thread_pool_to_notify_->NotifyDependenciesCompleted(this);
  }
}

void Task::AddDependentTask(Task* dependent_task) {
  std::lock_guard<std::mutex> lock(mutex_);
  if (state_ == COMPLETED) {
    dependent_task->OnDependenyCompleted();
    return;
  }
  CSM_CHECK(dependent_tasks_.insert(dependent_task).second);
}

void Task::OnDependenyCompleted() {
  std::lock_guard<std::mutex> lock(mutex_);
  CSM_CHECK(state_ == NEW || state_ == DISPATCHED);
  --uncompleted_dependencies_;
  if (uncompleted_dependencies_ == 0 && state_ == DISPATCHED) {
    state_ = DEPENDENCIES_COMPLETED;
    thread_pool_to_notify_->NotifyDependenciesCompleted(this);
  }
}

void Task::Execute() {
  {
    std::lock_guard<std::mutex> lock(mutex_);
    CSM_CHECK(state_ == DEPENDENCIES_COMPLETED);
    state_ = RUNNING;
  }
  if (work_item_) work_item_();
  std::lock_guard<std::mutex> lock(mutex_);
  state_ = COMPLETED;
  for (Task* dependent : dependent_tasks_) dependent->OnDependenyCompleted();
}

// ---- ThreadPool (cartographer/common/thread_pool.cc:32-104) --------------------
void ThreadPoolInterface::Execute(Task* task) { task->Execute(); }
void ThreadPoolInterface::SetThreadPool(Task* task) { task->SetThreadPool(this); }

ThreadPool::ThreadPool(int num_threads) {
  CSM_CHECK(num_threads > 0);
  std::lock_guard<std::mutex> lock(mutex_);
  for (int i = 0; i != num_threads; ++i) pool_.emplace_back([this]() { DoWork(); });
}

ThreadPool::~ThreadPool() {
  {
    std::lock_guard<std::mutex> lock(mutex_);
    CSM_CHECK(running_);
    running_ = false;
  }
  wake_.notify_all();
  for (std::thread& t : pool_) t.join();
}

void ThreadPool::NotifyDependenciesCompleted(Task* task) {
  {
    std::lock_guard<std::mutex> lock(mutex_);
    auto it = tasks_not_ready_.find(task);
    CSM_CHECK(it != tasks_not_ready_.end());
    task_queue_.push_back(it->second);
    tasks_not_ready_.erase(it);
  }
  wake_.notify_one();
}

std::weak_ptr<Task> ThreadPool::Schedule(std::unique_ptr<Task> task) {
  std::shared_ptr<Task> shared_task;
  {
    std::lock_guard<std::mutex> lock(mutex_);
    auto insert_result = tasks_not_ready_.insert(std::make_pair(task.get(), std::move(task)));
    CSM_CHECK(insert_result.second);
    shared_task = insert_result.first->second;
  }
  SetThreadPool(shared_task.get());
  return shared_task;
}

void ThreadPool::DoWork() {
  // Background work runs at low priority, like the reference (thread_pool.cc:74-79).
  if (nice(10) == -1) {
    // ignore: priority is best effort
  }
  for (;;) {
    std::shared_ptr<Task> task;
    {
      std::unique_lock<std::mutex> lock(mutex_);
      wake_.wait(lock, [this]() { return !task_queue_.empty() || !running_; });
      if (!task_queue_.empty()) {
        task = std::move(task_queue_.front());
        task_queue_.pop_front();
      } else if (!running_) {
        return;
      }
    }
    CSM_CHECK(task != nullptr);
    Execute(task.get());
  }
}

}  // namespace common
}  // namespace cartographer
