// Peer lists taken from a hosting platform's environment (no kungfu-run).
#pragma once
#include <string>

#include "core/plan.hpp"

namespace kf {

// Huawei ModelArts: DLS_TASK_INDEX, DLS_TASK_NUMBER, BATCH_CUSTOM<i>_HOSTS.
// Returns 1 and fills workers / self_task when the variables describe a
// cluster, 0 when the process is not running under ModelArts, -1 (+ *err)
// when the variables are present but unusable.
int modelarts_cluster(PeerList *workers, int *self_task, std::string *err);

}  // namespace kf
