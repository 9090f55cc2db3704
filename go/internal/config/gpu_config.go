// Source-only: not compiled in this repository's build image (no Go toolchain). See INTEGRATION.md.

package config

// GPUConfig is the `gpu:` section a CUDA box needs and the reference's config has no place for: its
// WorkerConfig is only the pool-facing worker name (config.go:225-229) and startMinerWorkers gives every
// SHA256d device a worker with default settings (setup.go:59-77). Added as a field `GPU GPUConfig
// `yaml:"gpu"`` of Config; the zero value keeps today's behaviour (every device, library defaults).
type GPUConfig struct {
	// Devices is an allow-list of hal.Identity.ID strings ("gpu-cuda0", ...). Empty = every CUDA device.
	// An ID that is not detected at start-up is a configuration error, not a silently idle GPU.
	Devices []string `yaml:"devices"`
	// Defaults apply to every CUDA device; PerDevice overrides them by device ID.
	Defaults  GPUDeviceConfig            `yaml:"defaults"`
	PerDevice map[string]GPUDeviceConfig `yaml:"per_device"`
}

// GPUDeviceConfig maps one to one onto otd_worker_config_t (include/otedama_b200.h).
type GPUDeviceConfig struct {
	RingCapacity  uint32 `yaml:"ring_capacity"`  // device share ring entries per launch; 0 = 4096
	QueueCapacity uint32 `yaml:"queue_capacity"` // host share queue; 0 = 65536
	LaunchLog2    uint32 `yaml:"launch_log2"`    // candidates per kernel launch = 2^n, 16..32; 0 = 32
	KernelVariant *int32 `yaml:"kernel_variant"` // nil = library default
}

// For returns the settings of one device: its own entry if present, else the defaults.
func (g GPUConfig) For(deviceID string) GPUDeviceConfig {
	if c, ok := g.PerDevice[deviceID]; ok {
		return c
	}
	return g.Defaults
}

// Allowed reports whether the allow-list lets deviceID mine.
func (g GPUConfig) Allowed(deviceID string) bool {
	if len(g.Devices) == 0 {
		return true
	}
	for _, id := range g.Devices {
		if id == deviceID {
			return true
		}
	}
	return false
}
