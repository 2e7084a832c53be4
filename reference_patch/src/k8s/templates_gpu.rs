//! UNVERIFIED (never compiled) — the two conditional blocks `--gpus` adds to the StatefulSet (templates.rs:9-57).
//!
//! The reference fills `STATEFUL_SET_TEMPLATE` by placeholder substitution (templates.rs:59-72).  The patch adds two
//! placeholders that expand to NOTHING when `gpus == 0`, so `h2ok deploy --cluster_size 3` renders byte for byte what it
//! renders today (pinned by tests/test_config1_plumbing.py through the Python restatement of this file).
//!
//!   resources.limits / resources.requests  gain   nvidia.com/gpu: '<gpus>'   (templates.rs:41-47)
//!   the pod spec                           gains  nodeSelector: nvidia.com/gpu.product: NVIDIA-B200   (templates.rs:26-27)
//!
//! With 8 GPUs per pod on one 8-GPU node, `replicas` x `gpus` must not exceed what the node has; the scheduler leaves
//! the surplus pods Pending, which `deploy` already reports through the readiness probe timing out.

/// Replaces `<gpu-resources>` (inside both `limits:` and `requests:`; indentation of 14 spaces like `cpu:`).
pub fn gpu_resources_block(gpus: u32) -> String {
    if gpus == 0 {
        return String::new();
    }
    format!("\n              nvidia.com/gpu: '{}'", gpus)
}

/// Replaces `<gpu-node-selector>` (sibling of `containers:`, indentation of 6 spaces).
pub fn gpu_node_selector_block(gpus: u32) -> String {
    if gpus == 0 {
        return String::new();
    }
    "\n      nodeSelector:\n        nvidia.com/gpu.product: NVIDIA-B200".to_string()
}
