/*
Command goptuna-suggestion (B200 build): the same listener as the reference's
cmd/suggestion/goptuna/v1beta1/main.go:28-57 -- plaintext gRPC on 0.0.0.0:6789, api.v1.beta1.Suggestion plus
grpc.health.v1.Health answering SERVING -- in front of the cgo service of
pkg/suggestion/v1beta1/goptuna/service.go in this tree.

Two differences from the reference, both supersets:
  - grpc.MaxRecvMsgSize is raised: the controller resends every trial on every call
    (suggestionclient.go:111-116), and a 10 000-trial x 50-parameter request is about 18 MB, above the 4 MiB
    default the reference server runs with (main.go:48);
  - KATIB_B200_DEVICES selects the GPUs (default "0").
*/
package main

import (
	"context"
	"net"
	"os"
	"os/signal"
	"syscall"

	health_pb "github.com/kubeflow/katib/pkg/apis/manager/health"
	api_v1_beta1 "github.com/kubeflow/katib/pkg/apis/manager/v1beta1"
	suggestion "github.com/kubeflow/katib/pkg/suggestion/v1beta1/goptuna"
	"google.golang.org/grpc"
	"k8s.io/klog"
)

const (
	address        = "0.0.0.0:6789"
	maxRecvMsgSize = 256 << 20
)

type healthService struct{}

// Check ignores the service name, like the reference (main.go:37-41); the composer's probes ask for
// manager.v1beta1.Suggestion (consts/const.go:170).
func (s *healthService) Check(ctx context.Context, in *health_pb.HealthCheckRequest) (*health_pb.HealthCheckResponse, error) {
	return &health_pb.HealthCheckResponse{Status: health_pb.HealthCheckResponse_SERVING}, nil
}

func main() {
	l, err := net.Listen("tcp", address)
	if err != nil {
		klog.Fatalf("Failed to listen: %v", err)
	}
	svc := suggestion.NewSuggestionService()
	srv := grpc.NewServer(grpc.MaxRecvMsgSize(maxRecvMsgSize))
	api_v1_beta1.RegisterSuggestionServer(srv, svc)
	health_pb.RegisterHealthServer(srv, &healthService{})

	// release the GPU state on SIGTERM (pod deletion) before the process goes away
	sig := make(chan os.Signal, 1)
	signal.Notify(sig, syscall.SIGTERM, syscall.SIGINT)
	go func() {
		<-sig
		srv.GracefulStop()
		svc.Close()
	}()

	klog.Infof("Start Goptuna suggestion service (B200 engine, devices %q): %s", os.Getenv(suggestion.DevicesEnv), address)
	if err = srv.Serve(l); err != nil {
		klog.Fatalf("Failed to serve: %v", err)
	}
}
