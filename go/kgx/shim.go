// Package kgx binds libkgx.so (include/kgx.h) for the exporter's Go host.
//
// NOT COMPILED ON THE BUILD BOX: there is no Go toolchain there (`go: command
// not found`).  It is the reference-side binding a maintainer adds next to
// main.go; see INTEGRATION.md.
package kgx

/*
#cgo CFLAGS: -I${SRCDIR}/../../include
#cgo LDFLAGS: -L${SRCDIR}/../../kubernetes-gpu-exporter_b200/csrc -lkgx
#include <stdlib.h>
#include "kgx.h"
*/
import "C"

import (
	"fmt"
	"io"
	"runtime"
	"unsafe"
)

type Mode int

const (
	ModeRIntended Mode = C.KGX_MODE_R_INTENDED // main.go's families, PID value equality
	ModeRLiteral  Mode = C.KGX_MODE_R_LITERAL  // main.go:144-145 exactly as written
	ModeX         Mode = C.KGX_MODE_X          // (pod, namespace, container, gpu) reductions
)

type Ctx struct{ p *C.kgx_ctx }

func (c *Ctx) err(rc C.int) error {
	if rc == C.KGX_OK {
		return nil
	}
	return fmt.Errorf("kgx: %s: %s", C.GoString(C.kgx_strerror(rc)), C.GoString(C.kgx_last_error(c.p)))
}

func New(device int, mode Mode) (*Ctx, error) {
	opts := C.kgx_opts{abi_version: C.KGX_ABI_VERSION, mode: C.int32_t(mode)}
	c := &Ctx{}
	if rc := C.kgx_create(C.int(device), &opts, &c.p); rc != C.KGX_OK {
		return nil, fmt.Errorf("kgx_create: %s (a CUDA device is required; there is no CPU fallback)", C.GoString(C.kgx_strerror(rc)))
	}
	runtime.SetFinalizer(c, func(c *Ctx) { c.Close() })
	return c, nil
}

func (c *Ctx) Close() {
	if c.p != nil {
		C.kgx_destroy(c.p)
		c.p = nil
	}
}

// Table is the flat form of `podPIDMap map[string][]string` (main.go:84).
// The library copies everything before returning, so Go memory may be passed
// directly (no Go pointer is retained — the cgo rule).
type Table struct {
	Nodes, GPUsPerNode uint32
	PodkeyNode         []uint32
	Pod, Namespace, Container []string // Namespace/Container: X mode only
	EntryPodkey, EntryPID     []uint32
	GPUUUID                   []string // X mode only, index node*G+gpu
}

func pack(ss []string) (blob []byte, off []uint64) {
	off = make([]uint64, len(ss)+1)
	for i, s := range ss {
		blob = append(blob, s...)
		off[i+1] = uint64(len(blob))
	}
	blob = append(blob, 0)
	return
}

func u32p(s []uint32) *C.uint32_t {
	if len(s) == 0 {
		return nil
	}
	return (*C.uint32_t)(unsafe.Pointer(&s[0]))
}
func u64p(s []uint64) *C.uint64_t {
	if len(s) == 0 {
		return nil
	}
	return (*C.uint64_t)(unsafe.Pointer(&s[0]))
}

func (c *Ctx) SetTable(t *Table) error {
	var pin runtime.Pinner
	defer pin.Unpin()
	ct := C.kgx_table{n_nodes: C.uint32_t(t.Nodes), gpus_per_node: C.uint32_t(t.GPUsPerNode),
		n_podkeys: C.uint64_t(len(t.PodkeyNode)), n_entries: C.uint64_t(len(t.EntryPID))}
	set := func(ss []string, blob **C.char, off **C.uint64_t) {
		if ss == nil {
			return
		}
		b, o := pack(ss)
		pin.Pin(&b[0])
		pin.Pin(&o[0])
		*blob, *off = (*C.char)(unsafe.Pointer(&b[0])), u64p(o)
	}
	set(t.Pod, &ct.pod_blob, &ct.pod_off)
	set(t.Namespace, &ct.namespace_blob, &ct.namespace_off)
	set(t.Container, &ct.container_blob, &ct.container_off)
	set(t.GPUUUID, &ct.gpu_uuid_blob, &ct.gpu_uuid_off)
	for _, s := range [][]uint32{t.PodkeyNode, t.EntryPodkey, t.EntryPID} {
		if len(s) > 0 {
			pin.Pin(&s[0])
		}
	}
	ct.podkey_node, ct.entry_podkey, ct.entry_pid = u32p(t.PodkeyNode), u32p(t.EntryPodkey), u32p(t.EntryPID)
	return c.err(C.kgx_set_table(c.p, &ct))
}

// Cols is one poll cycle's samples in column form (main.go:123-138).
type Cols struct {
	PID, GPU, Node, SMUtil, MemUtil []uint32
	Used, TsUs, NodeOff, GPUTotal   []uint64
}

func (c *Ctx) Ingest(s *Cols) error {
	var pin runtime.Pinner
	defer pin.Unpin()
	for _, a := range [][]uint32{s.PID, s.GPU, s.Node, s.SMUtil, s.MemUtil} {
		if len(a) > 0 {
			pin.Pin(&a[0])
		}
	}
	for _, a := range [][]uint64{s.Used, s.TsUs, s.NodeOff, s.GPUTotal} {
		if len(a) > 0 {
			pin.Pin(&a[0])
		}
	}
	cc := C.kgx_cols{n: C.uint64_t(len(s.PID)), pid: u32p(s.PID), gpu: u32p(s.GPU), used: u64p(s.Used), node: u32p(s.Node),
		sm_util: u32p(s.SMUtil), mem_util: u32p(s.MemUtil), ts_us: u64p(s.TsUs), node_off: u64p(s.NodeOff), gpu_total: u64p(s.GPUTotal)}
	return c.err(C.kgx_ingest(c.p, &cc))
}

// IngestNvml feeds the R modes straight from go-nvml: infos is every device's
// GetComputeRunningProcesses result concatenated in the loop order of main.go:123
// (nvml.ProcessInfo has the C layout of nvmlProcessInfo_t), devOff the row range of
// each device.  IngestNvmlX does the same for X mode with the utilisation samples
// of nvml.Device.GetProcessUtilization beside them.
func (c *Ctx) IngestNvml(infos []C.kgx_nvml_process_info, devOff, gpuTotal []uint64) error {
	var pin runtime.Pinner
	defer pin.Unpin()
	var ip *C.kgx_nvml_process_info
	if len(infos) > 0 {
		pin.Pin(&infos[0])
		ip = &infos[0]
	}
	pin.Pin(&devOff[0])
	pin.Pin(&gpuTotal[0])
	return c.err(C.kgx_ingest_nvml(c.p, ip, u64p(devOff), u64p(gpuTotal)))
}

func (c *Ctx) IngestNvmlX(infos []C.kgx_nvml_process_info, devOff []uint64, utils []C.kgx_nvml_process_utilization_sample,
	utilOff, gpuTotal []uint64) error {
	var pin runtime.Pinner
	defer pin.Unpin()
	var ip *C.kgx_nvml_process_info
	var up *C.kgx_nvml_process_utilization_sample
	if len(infos) > 0 {
		pin.Pin(&infos[0])
		ip = &infos[0]
	}
	if len(utils) > 0 {
		pin.Pin(&utils[0])
		up = &utils[0]
	}
	pin.Pin(&devOff[0])
	pin.Pin(&utilOff[0])
	pin.Pin(&gpuTotal[0])
	return c.err(C.kgx_ingest_nvml_x(c.p, ip, u64p(devOff), up, u64p(utilOff), u64p(gpuTotal)))
}

func (c *Ctx) ResetWindow() error { return c.err(C.kgx_reset_window(c.p)) }

// Expire drops the X-mode series whose newest sample is older than the cutoff
// (microseconds, the clock of Cols.TsUs) and returns how many exported series
// went.  R modes refuse: the reference never deletes a GaugeVec child.
func (c *Ctx) Expire(olderThanUs uint64) (uint64, error) {
	var n C.uint64_t
	rc := C.kgx_expire(c.p, C.uint64_t(olderThanUs), &n)
	return uint64(n), c.err(rc)
}

// Render returns the text exposition; buf is reused when large enough.
func (c *Ctx) Render(buf []byte) ([]byte, error) {
	for {
		var n C.size_t
		var p *C.uint8_t
		if cap(buf) > 0 {
			p = (*C.uint8_t)(unsafe.Pointer(&buf[:1][0]))
		}
		rc := C.kgx_render(c.p, p, C.size_t(cap(buf)), &n)
		if rc == C.KGX_E_NOSPACE {
			buf = make([]byte, 0, int(n)+int(n)/8)
			continue
		}
		if err := c.err(rc); err != nil {
			return nil, err
		}
		return buf[:int(n)], nil
	}
}

// RenderDevice renders into device memory and returns the size of the body.
func (c *Ctx) RenderDevice() (int, error) {
	var n C.size_t
	rc := C.kgx_render_device(c.p, &n)
	return int(n), c.err(rc)
}

// WriteBodyTo renders on the device and writes the body to w chunk by chunk
// (len(chunk) bytes at a time, chunk reused): host memory stays at one chunk
// whatever the size of the body, and the first bytes leave one render and one
// chunk copy after the call.  The caller serialises scrapes on this Ctx.
func (c *Ctx) WriteBodyTo(w io.Writer, chunk []byte) (int64, error) {
	if len(chunk) == 0 {
		chunk = make([]byte, 1<<20)
	}
	if _, err := c.RenderDevice(); err != nil {
		return 0, err
	}
	var at int64
	for {
		var n C.size_t
		rc := C.kgx_body_read(c.p, C.uint64_t(at), (*C.uint8_t)(unsafe.Pointer(&chunk[0])), C.size_t(len(chunk)), &n)
		if err := c.err(rc); err != nil {
			return at, err
		}
		if n == 0 {
			return at, nil
		}
		if _, err := w.Write(chunk[:int(n)]); err != nil {
			return at, err
		}
		at += int64(n)
	}
}

type FamilyInfo struct {
	Name, Help string
	LabelNames []string
}

func (c *Ctx) Families() []FamilyInfo {
	n := int(C.kgx_family_count(c.p))
	out := make([]FamilyInfo, n)
	for f := 0; f < n; f++ {
		var name, help *C.char
		var names **C.char
		var nl C.int
		C.kgx_family_info(c.p, C.int(f), &name, &help, &names, &nl)
		out[f] = FamilyInfo{Name: C.GoString(name), Help: C.GoString(help)}
		for _, p := range unsafe.Slice(names, int(nl)) {
			out[f].LabelNames = append(out[f].LabelNames, C.GoString(p))
		}
	}
	return out
}

// ReadValues returns the values and slot ids of one family in exposition order.
func (c *Ctx) ReadValues(family int) ([]float64, []uint32, error) {
	var n C.size_t
	C.kgx_read_values(c.p, C.int(family), nil, nil, 0, &n)
	vals, slots := make([]float64, int(n)), make([]uint32, int(n))
	if n == 0 {
		return vals, slots, nil
	}
	rc := C.kgx_read_values(c.p, C.int(family), (*C.double)(unsafe.Pointer(&vals[0])), u32p(slots), n, &n)
	return vals[:int(n)], slots[:int(n)], c.err(rc)
}

// ReadLabels returns the pre-escaped `{name="value",...}` text of every series that has a value, in exposition order
// (entry i belongs to value i of every family), with one call into the library instead of one per series.
func (c *Ctx) ReadLabels() ([]string, error) {
	var n, nb C.size_t
	if err := c.err(C.kgx_read_labels(c.p, nil, 0, nil, 0, &n, &nb)); err != nil {
		return nil, err
	}
	blob := make([]byte, int(nb)+1)
	off := make([]uint64, int(n)+1)
	rc := C.kgx_read_labels(c.p, (*C.char)(unsafe.Pointer(&blob[0])), nb, u64p(off), C.size_t(len(off)), &n, &nb)
	if err := c.err(rc); err != nil {
		return nil, err
	}
	out := make([]string, int(n))
	for i := range out {
		out[i] = string(blob[off[i]:off[i+1]])
	}
	return out, nil
}

// SlotLabels returns the pre-escaped `{name="value",...}` text of a slot.
func (c *Ctx) SlotLabels(slot uint32) string {
	buf := make([]byte, 1024)
	var n C.size_t
	C.kgx_slot_labels(c.p, C.uint32_t(slot), (*C.char)(unsafe.Pointer(&buf[0])), C.size_t(len(buf)), &n)
	if int(n) > len(buf) {
		buf = make([]byte, int(n))
		C.kgx_slot_labels(c.p, C.uint32_t(slot), (*C.char)(unsafe.Pointer(&buf[0])), C.size_t(len(buf)), &n)
	}
	return string(buf[:int(n)])
}
