/*
Package suggestion_goptuna_v1beta1 is the B200 replacement for Katib's goptuna suggestion service
(reference: pkg/suggestion/v1beta1/goptuna/service.go:47-215, converter.go, sample.go).

It keeps the reference's Go surface -- NewSuggestionService, the api.v1.beta1.SuggestionServer methods
GetSuggestions and ValidateAlgorithmSettings -- and forwards the request bytes over cgo to libkatib_b200.so
(include/katib_b200.h: kb_service_*), which restates service.go / converter.go / sample.go above the CUDA kernels
and answers with the serialized reply and the gRPC status code the reference returns at the same spot.

This file drops into the reference tree in place of service.go, converter.go and sample.go of that package; the
generated api.pb.go and the controller stay untouched.  It is built with CGO_ENABLED=1 against the header and the
library (see the Dockerfile next to main.go).  There is no Go toolchain in the build image of this repository, so the
file is kept small and is checked by reading, not by a compiler.
*/
package suggestion_goptuna_v1beta1

/*
#cgo CFLAGS: -I/usr/local/include
#cgo LDFLAGS: -L/usr/local/lib -lkatib_b200
#include <stdlib.h>
#include "katib_b200.h"
*/
import "C"

import (
	"context"
	"fmt"
	"os"
	"strconv"
	"strings"
	"sync"
	"unsafe"

	"google.golang.org/grpc/codes"
	"google.golang.org/grpc/status"
	"google.golang.org/protobuf/proto"

	api_v1_beta1 "github.com/kubeflow/katib/pkg/apis/manager/v1beta1"
)

// DevicesEnv lists the CUDA devices one service instance drives, e.g. "0" or "0,1,2,3,4,5,6,7".
// Unset = device 0.  With more than one device the candidate set of a large TPE request is sharded over them and the
// per-GPU winners are merged inside the library (kb_service_create_multi).
const DevicesEnv = "KATIB_B200_DEVICES"

// SuggestionService replaces the struct at service.go:47-52: the study, the search space and the trial mapping all
// live behind the kb_service handle.
type SuggestionService struct {
	api_v1_beta1.UnimplementedSuggestionServer
	mu  sync.Mutex // the reference serialises init and syncTrials (service.go:97-98, 156-157); one handle, one caller
	svc *C.kb_service
	err error // why the engine could not be created (no B200, no library): every call then answers codes.Internal
}

// NewSuggestionService keeps the reference's constructor name (service.go:39-45).
func NewSuggestionService() *SuggestionService {
	s := &SuggestionService{}
	devices, err := parseDevices(os.Getenv(DevicesEnv))
	if err != nil {
		s.err = err
		return s
	}
	var h *C.kb_service
	var rc C.int
	if len(devices) == 1 {
		rc = C.kb_service_create(C.int(devices[0]), &h)
	} else {
		cdev := make([]C.int, len(devices))
		for i, d := range devices {
			cdev[i] = C.int(d)
		}
		rc = C.kb_service_create_multi(C.int(len(cdev)), &cdev[0], &h)
	}
	if rc != 0 {
		s.err = fmt.Errorf("katib_b200: cannot create the engine on devices %v: %s", devices, C.GoString(C.kb_last_error(nil)))
		return s
	}
	s.svc = h
	return s
}

func parseDevices(v string) ([]int, error) {
	if strings.TrimSpace(v) == "" {
		return []int{0}, nil
	}
	var out []int
	for _, f := range strings.Split(v, ",") {
		d, err := strconv.Atoi(strings.TrimSpace(f))
		if err != nil || d < 0 {
			return nil, fmt.Errorf("%s=%q: not a list of device indices", DevicesEnv, v)
		}
		out = append(out, d)
	}
	return out, nil
}

// Close releases the GPU state (the reference has nothing to release).
func (s *SuggestionService) Close() {
	s.mu.Lock()
	defer s.mu.Unlock()
	if s.svc != nil {
		C.kb_service_destroy(s.svc)
		s.svc = nil
	}
}

// bytesPtr gives C a pointer to the message bytes; an empty message still needs a valid pointer for cgo.
func bytesPtr(b []byte) (unsafe.Pointer, C.size_t) {
	if len(b) == 0 {
		var zero [1]byte
		return unsafe.Pointer(&zero[0]), 0
	}
	return unsafe.Pointer(&b[0]), C.size_t(len(b))
}

// GetSuggestions: service.go:54-93.  Every failure is codes.Internal with the reference's message text.
func (s *SuggestionService) GetSuggestions(ctx context.Context, req *api_v1_beta1.GetSuggestionsRequest) (*api_v1_beta1.GetSuggestionsReply, error) {
	if s.err != nil {
		return nil, status.Error(codes.Internal, s.err.Error())
	}
	in, err := proto.Marshal(req)
	if err != nil {
		return nil, status.Error(codes.Internal, err.Error())
	}
	s.mu.Lock()
	defer s.mu.Unlock()
	p, n := bytesPtr(in)
	var out unsafe.Pointer
	var outLen C.size_t
	code := C.kb_service_get_suggestions(s.svc, p, n, &out, &outLen)
	if code != 0 {
		return nil, status.Error(codes.Code(code), C.GoString(C.kb_service_last_message(s.svc)))
	}
	reply := &api_v1_beta1.GetSuggestionsReply{}
	// the reply bytes belong to the handle until its next call: copy them out while the lock is held
	if err := proto.Unmarshal(C.GoBytes(out, C.int(outLen)), reply); err != nil {
		return nil, status.Error(codes.Internal, err.Error())
	}
	return reply, nil
}

// ValidateAlgorithmSettings: service.go:173-211 (InvalidArgument / Internal exactly as there; needs no GPU).
func (s *SuggestionService) ValidateAlgorithmSettings(ctx context.Context, req *api_v1_beta1.ValidateAlgorithmSettingsRequest) (*api_v1_beta1.ValidateAlgorithmSettingsReply, error) {
	if s.err != nil {
		return nil, status.Error(codes.Internal, s.err.Error())
	}
	s.mu.Lock()
	defer s.mu.Unlock()
	var code C.int
	if req == nil {
		code = C.kb_service_validate_algorithm_settings(s.svc, nil, 0) // the nil request of service.go:177-179
	} else {
		in, err := proto.Marshal(req)
		if err != nil {
			return nil, status.Error(codes.Internal, err.Error())
		}
		p, n := bytesPtr(in)
		code = C.kb_service_validate_algorithm_settings(s.svc, p, n)
	}
	if code != 0 {
		return nil, status.Error(codes.Code(code), C.GoString(C.kb_service_last_message(s.svc)))
	}
	return &api_v1_beta1.ValidateAlgorithmSettingsReply{}, nil
}

// compile-time check, as service.go:213-215
var _ api_v1_beta1.SuggestionServer = &SuggestionService{}
