// Source-only: not compiled in this repository's build image (no Go toolchain). See INTEGRATION.md.
//go:build cgo && cuda

package hal

/*
#cgo LDFLAGS: -lotedama_b200
#include "otedama_b200.h"
*/
import "C"

import (
	"context"
	"fmt"
	"runtime"
)

type cudaDevice struct {
	index int
	id    Identity
	caps  Capabilities
}

func (d *cudaDevice) Identity() Identity             { return d.id }
func (d *cudaDevice) Capabilities() Capabilities     { return d.caps }
func (d *cudaDevice) Shutdown(context.Context) error { return nil } // idempotent; workers own device state
func (d *cudaDevice) Index() int                     { return d.index }

// CUDAIndexed is implemented by devices of this driver: the engine needs the CUDA ordinal to open a worker.
type CUDAIndexed interface{ Index() int }

// CUDADriver implements hal.Driver (internal/hal/device.go:182-198). It is registered next to
// GPULinuxDriver in engine.detectDevices (setup.go:29-36); the sysfs driver keeps reporting SHA256d:false
// for the same boards (gpu_linux.go:20-31), this one reports what the library's kernels can do.
type CUDADriver struct{}

func (CUDADriver) Name() string { return "cuda" }

func (CUDADriver) Enumerate(ctx context.Context) ([]Device, error) {
	// otd_last_error() is thread-local: keep the failing call and the read of its message on one OS thread
	runtime.LockOSThread()
	defer runtime.UnlockOSThread()
	n := int(C.otd_device_count())
	if n < 0 {
		return nil, fmt.Errorf("hal/cuda: %s", C.GoString(C.otd_last_error()))
	}
	devs := make([]Device, 0, n)
	for i := 0; i < n; i++ {
		if err := ctx.Err(); err != nil {
			return devs, err // partial result + ctx.Err(), as the interface asks
		}
		var info C.otd_device_info_t
		if C.otd_device_info(C.int(i), &info) != 0 {
			continue // one bad board does not hide the others (registry.go:180-190 logs per driver)
		}
		devs = append(devs, &cudaDevice{
			index: i,
			id: Identity{ID: C.GoString(&info.id[0]), Family: FamilyGPU,
				Vendor: C.GoString(&info.vendor[0]), Model: C.GoString(&info.model[0])},
			caps: Capabilities{SHA256d: info.cap_sha256d != 0, GeneralCompute: info.cap_general_compute != 0},
		})
	}
	return devs, nil
}
