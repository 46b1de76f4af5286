;;;; fusion.lisp — IR rewrite hook + CUDA-graph replay of a compiled iseq + data-parallel helpers.
;;;;
;;;; `on-finalizing-compiling` is called once per `build` per device with the forward / backward
;;;; InstructionSeq (`source/vm/nodes/defnode.lisp:26-57`, `vm/compile.lisp:237-238`; working
;;;; precedent: `source/backends/aten/tensor.lisp:105-141`).  Three rewrites are installed:
;;;;   1. MoveTensorNode immediately followed by an in-place {Add,Sub,Mul,Div}Node on the copy
;;;;      (`!add` = `!copy` + `A+=B`, `base-impl/arithmetic.lisp:150-152`) -> one wf2cu_ew_binary3 launch.
;;;;   2. MatMulNode followed by the in-place AddNode of LinearLayer's `(!flexible bias)` operand
;;;;      (`nn/regression.lisp:14-18`) -> one wf2cu_gemm_ex with the bias in the epilogue.
;;;;   3. Im2ColNode + Permute-Node + the `->contiguous` MoveTensorNode of an unpadded `unfold`
;;;;      (`nn/im2col.lisp:15-54`) -> one wf2cu_unfold.
;;;; Two composites are offered as drop-in functions instead of pattern matches (they are plain functions in the
;;;; reference, not nodes): `cuda-layer-norm` (the GPT-2 example's LayerNorm-Revisit) and `cuda-softmax` (`!softmax`,
;;;; optionally with the `(!div w n)` in front of it), forward only.
;;;; NOT rewritten here yet — bound in cffi.lisp and driven through the same C ABI by the Python host
;;;; (cl-waffe2_b200/nn.py, conv.py), where they are tested: the `!relu` / relu-backward-mask forms of wf2cu_gemm_ex,
;;;; wf2cu_sce_sum_forward/backward, wf2cu_fold, wf2cu_pool2d(_backward), wf2cu_ew_scalar for non-constant scalars
;;;; (a constant scalar operand is a one-element CUDATensor made at build time and viewed with (:broadcast n): no
;;;; per-step transfer; a ScalarTensor goes through Scalar->MatNode's 4-byte `(setf vref)`).
;;;; Everything else is left as generated; each remaining instruction is already one kernel.
(in-package :cl-waffe2/backends.cuda)

(defun arith-op-code (node)
  (typecase node
    (AddNode +add+) (SubNode +sub+) (MulNode +mul+) (DivNode +div+) (t nil)))

(defun cuda-inst-p (inst)
  (every #'(lambda (x) (typep x 'CUDATensor)) (wf/vm:wfop-args inst)))

(defun fuse-move+arith (iseq)
  "copy <- MoveTensor(copy, x) ; copy <- Arith(copy, y)   ==>   copy <- binary3(op, copy, x, y)"
  (loop with out = nil
	while iseq
	do (let* ((i0 (pop iseq))
		  (i1 (car iseq)))
	     (if (and i1
		      (typep (wf/vm:wfop-node i0) 'MoveTensorNode)
		      (arith-op-code (wf/vm:wfop-node i1))
		      (cuda-inst-p i0) (cuda-inst-p i1)
		      (not (movetensor-ignore-me (wf/vm:wfop-node i0)))
		      (null (wf/vm:wfop-sv4bw i1))
		      ;; the arith node works in place on the Move's destination
		      (eq (first (wf/vm:wfop-out-to i0)) (first (wf/vm:wfop-args i1))))
		 (let ((code (arith-op-code (wf/vm:wfop-node i1))))
		   (pop iseq)
		   ;; keep i1's bookkeeping (out-to, node, self), replace its closure and arguments
		   (setf (wf/vm:wfop-args i1) (list (first (wf/vm:wfop-args i0))
						    (second (wf/vm:wfop-args i0))
						    (second (wf/vm:wfop-args i1)))
			 (wf/vm:wfop-op i1)
			 #'(lambda (dst x y)
			     (with-lowered-views ((dv dst) (xv x) (yv y))
			       (ew-binary3 code dv xv yv))
			     dst))
		   (push i1 out))
		 (push i0 out)))
	finally (return (nreverse out))))

(defun bias-operand-p (tensor n)
  "T when TENSOR is a rank-2 view whose row axis is (:broadcast _) over a contiguous N-element vector:
what `(!flexible bias)` becomes inside `step-linear` (`nn/regression.lisp:14-18`)."
  (and (typep tensor 'CUDATensor)
       (= (dims tensor) 2)
       (subscript-broadcast (first (tensor-view tensor)))
       (= (observe (second (shape tensor))) n)
       (eql (observe (second (tensor-stride tensor))) 1)))

(defun fuse-matmul+bias (iseq)
  "out <- MatMul(a, b, out) ; out <- Add(out, bias*)   ==>   out <- gemm_ex(a, b, out, bias)"
  (loop with result = nil
	while iseq
	do (let* ((i0 (pop iseq))
		  (i1 (car iseq)))
	     (if (and i1
		      (typep (wf/vm:wfop-node i0) 'MatMulNode)
		      (typep (wf/vm:wfop-node i1) 'AddNode)
		      (cuda-inst-p i0) (cuda-inst-p i1)
		      (null (wf/vm:wfop-sv4bw i1))
		      (= (dims (third (wf/vm:wfop-args i0))) 2)
		      (eq (first (wf/vm:wfop-out-to i0)) (first (wf/vm:wfop-args i1)))
		      (bias-operand-p (second (wf/vm:wfop-args i1))
				      (observe (second (shape (third (wf/vm:wfop-args i0)))))))
		 (let ((node (wf/vm:wfop-node i0))
		       (bias (second (wf/vm:wfop-args i1))))
		   (pop iseq)
		   (setf (wf/vm:wfop-args i1) (append (wf/vm:wfop-args i0) (list bias))
			 (wf/vm:wfop-op i1)
			 #'(lambda (a b out bias)
			     (let* ((ta (trans-a? node)) (tb (trans-b? node))
				    (a* (if ta (read-untransposed a) a))
				    (b* (if tb (read-untransposed b) b))
				    (m (observe (first (shape out)))) (n (observe (second (shape out))))
				    (k (observe (car (last (shape a))))))
			       (with-lowered-views ((av a*) (bv b*) (cv out) (biasv bias))
				 (flet ((off (v) (foreign-slot-value v '(:struct wf2cu-view) 'offset))
					(ld  (v) (mem-aref (foreign-slot-pointer v '(:struct wf2cu-view) 'stride) :int64 0)))
				   (gemm-ex (if ta 1 0) (if tb 1 0) m n k
					    (device-vec-ptr (tensor-vec a*)) (off av) (ld av)
					    (device-vec-ptr (tensor-vec b*)) (off bv) (ld bv)
					    (device-vec-ptr (tensor-vec out)) (off cv) (ld cv)
					    (device-vec-ptr (tensor-vec bias)) (off biasv) 0
					    (null-pointer) 0 0))))
			     out))
		   (push i1 result))
		 (push i0 result)))
	finally (return (nreverse result))))

;; ---- unfold: Im2ColNode ; Permute-Node ; MoveTensor(->contiguous)  ==>  one wf2cu_unfold
;; `!im2col` (`nn/im2col.lisp:15-54`) always emits  col <- Im2Col(x, col) ; p <- Permute(col) ;
;; dense <- MoveTensor(dense, p)  (the `->contiguous` inside `!reshape`).  With padding (0 0) — MNIST-CNN,
;; the pooling layers — the three instructions become one launch that reads the image through its view
;; and writes the (N*Ho*Wo, C*kh*kw) rows directly.  With padding the chain is left as it is (the padded
;; image is produced by a Move into a window, and the node then skips trailing rows: three launches).
(defun unpadded-unfold-p (node)
  (with-slots (padding-h padding-w dilation-h dilation-w) node
    (and (eql padding-h 0) (eql padding-w 0) (eql dilation-h 1) (eql dilation-w 1))))

(defun fuse-unfold-chain (iseq)
  (loop with out = nil
	while iseq
	do (let* ((i0 (pop iseq))
		  (i1 (first iseq))
		  (i2 (second iseq)))
	     (if (and i1 i2
		      (typep (wf/vm:wfop-node i0) 'Im2ColNode)
		      (unpadded-unfold-p (wf/vm:wfop-node i0))
		      (typep (wf/vm:wfop-node i1) 'cl-waffe2/base-impl::Permute-Node)
		      (typep (wf/vm:wfop-node i2) 'MoveTensorNode)
		      (cuda-inst-p i0) (cuda-inst-p i2)
		      (null (wf/vm:wfop-sv4bw i0))
		      (eq (first (wf/vm:wfop-out-to i0)) (first (wf/vm:wfop-args i1)))
		      (eq (first (wf/vm:wfop-out-to i1)) (second (wf/vm:wfop-args i2))))
		 (let ((node (wf/vm:wfop-node i0))
		       (x    (first (wf/vm:wfop-args i0))))
		   (pop iseq) (pop iseq)
		   (setf (wf/vm:wfop-args i2) (list (first (wf/vm:wfop-args i2)) x)
			 (wf/vm:wfop-op i2)
			 #'(lambda (dense image)
			     (with-unfold-geom (g node (shape image))
			       (with-lowered-views ((xv image))
				 (unfold xv (device-vec-ptr (tensor-vec dense)) (tensor-initial-offset dense) g)))
			     dense))
		   (push i2 out))
		 (push i0 out)))
	finally (return (nreverse out))))

(defmethod on-finalizing-compiling ((device-name (eql 'CUDATensor)) iseq-fw iseq-bw)
  (if *cuda-available-p*
      (flet ((rewrite (iseq) (fuse-matmul+bias (fuse-move+arith (fuse-unfold-chain iseq)))))
	(values (rewrite iseq-fw) (and iseq-bw (rewrite iseq-bw))))
      (values iseq-fw iseq-bw)))

;; ---- LayerNorm over the last axis as ONE launch (csrc/softmax.cu `layernorm_kernel`).  The GPT-2 example
;; spells it as eleven nodes (`LayerNorm-Revisit`, examples/gpt-2/utils.lisp:114-127: two !mean, !sub, !expt,
;; !add eps, !sqrt, !div, !mul g, !add b); the kernel evaluates the same operations in the same order with
;; correctly rounded div/sqrt, so the result agrees with the node chain to the last few ulp.  Inference only
;; (the example never differentiates it): the backward is left to the node chain.
(define-op (CUDALayerNormNode (self eps)
	    :slots ((eps :initarg :eps :initform 1.0e-5 :reader eps-of))
	    :where (X[~ n] G[n] B[n] Out[~ n] -> Out[~ n])
	    :forward ((self x g b out)
		      (let* ((cols (car (last (shape x))))
			     (rows (/ (apply #'* (shape x)) cols)))
			(layernorm (device-vec-ptr (tensor-vec x)) (tensor-initial-offset x)
				   (device-vec-ptr (tensor-vec g)) (tensor-initial-offset g)
				   (device-vec-ptr (tensor-vec b)) (tensor-initial-offset b)
				   (device-vec-ptr (tensor-vec out)) (tensor-initial-offset out)
				   rows cols (coerce (eps-of self) 'single-float))
			out))
	    :backward ((self dout)
		       (declare (ignore dout))
		       (error "CUDALayerNormNode: forward only; differentiate LayerNorm-Revisit instead")))
  (setf (slot-value self 'eps) eps))

(defun cuda-layer-norm (x g b &key (eps 1.0e-5))
  "Drop-in for `LayerNorm-Revisit` on dense CUDATensors whose last axis is a multiple of 4 and <= 1024
(the library answers UNSUPPORTED otherwise, which `defwf2` signals as `cuda-backend-error`)."
  (call (CUDALayerNormNode eps) (->contiguous x) g b (make-input (shape x) nil :dtype (dtype x) :order (order x))))

;; ---- `!softmax` (`nn/activation.lisp:126-159`: five nodes and two disguised reductions) as ONE launch, optionally with the
;; `(!div w n)` that precedes it in the GPT-2 example folded in (`examples/gpt-2/utils.lisp:236-242`: PRE-DIVISOR = n; 0 =
;; none).  AXIS counts from the front like `!softmax :axis`.  Forward only, like CUDALayerNormNode.
(define-op (CUDASoftmaxNode (self axis avoid-overflow pre-divisor)
	    :slots ((axis :initform 1 :reader axis-of)
		    (avoid-overflow :initform t :reader avoid-overflow-of)
		    (pre-divisor :initform 0.0 :reader pre-divisor-of))
	    :where (X[~] Out[~] -> Out[~])
	    :forward ((self x out)
		      (with-lowered-views ((xv x) (ov out))
			(softmax-ex xv ov (axis-of self) (if (avoid-overflow-of self) 1 0)
				    (coerce (pre-divisor-of self) 'single-float)))
		      out)
	    :backward ((self dout)
		       (declare (ignore dout))
		       (error "CUDASoftmaxNode: forward only; differentiate !softmax instead")))
  (setf (slot-value self 'axis) axis
	(slot-value self 'avoid-overflow) avoid-overflow
	(slot-value self 'pre-divisor) pre-divisor))

(defun cuda-softmax (x &key (axis 1) (avoid-overflow t) (pre-divisor 0.0))
  "Drop-in for `(!softmax x :axis axis :avoid-overflow avoid-overflow)`; with PRE-DIVISOR n for `(!softmax (!div x n) ...)`."
  (call (CUDASoftmaxNode axis avoid-overflow pre-divisor) x (make-input (shape x) nil :dtype (dtype x) :order (order x))))

;; ---- independent instructions on a side stream: (with-cuda-branch (1) <launches>) ... (branch-join 1)
(defmacro with-cuda-branch ((b) &body body)
  "BODY's launches go to side stream B (1..3), ordered after everything issued so far; the main stream
waits for them at `(branch-join B)`.  Inside `with-cuda-graph` these become fork/join edges."
  `(progn (branch-begin ,b)
	  (unwind-protect (progn ,@body)
	    (branch-end ,b))))

;; ---- CUDA-graph replay (SURVEY §8f rank 1): after `set-adjustable-symbols` the iseq is a fixed
;; launch sequence; capture `(forward model)` once and replay it.
(defmacro with-cuda-graph ((exec-var) capture-form &body body)
  `(let ((,exec-var (with-foreign-object (out :pointer)
		      (graph-begin)
		      ,capture-form
		      (graph-end out)
		      (mem-ref out :pointer))))
     (unwind-protect (progn ,@body)
       (graph-destroy ,exec-var))))

;; ---- data-parallel MLP step (SURVEY §8e): one process per GPU; gradients live in one arena
;; (allocate parameters' grads as views of a single CUDATensor) so the bucket is one range.
(defun cuda-comm-init (rank world id-bytes)
  "ID-BYTES: the 128-byte NCCL unique id made by rank 0 (wf2cu_comm_unique_id) and shared by file/env."
  (with-foreign-object (id :uint8 128)
    (dotimes (i 128) (setf (mem-aref id :uint8 i) (aref id-bytes i)))
    (comm-init rank world id)))

(defun cuda-allreduce-gradients (grad-arena n-floats)
  "Sum the gradient arena over ranks on the library stream; the 1/world factor is applied by the
optimizer kernel (`grad_scale` of wf2cu_adam_step)."
  (allreduce-sum-f32 (device-vec-ptr (tensor-vec grad-arena)) (tensor-initial-offset grad-arena) n-floats))

;; ---- exchange + Adam as ONE kernel over NVLink peer memory (csrc/p2p.cu): reduce-scatter, Adam on this
;; rank's shard, parameter all-gather.  HANDLES-EXCHANGE is any host transport that all-gathers 192 bytes per
;; rank (file, socket, MPI): it receives this rank's (unsigned-byte 8) vector and returns a list of WORLD vectors.
(defun cuda-p2p-attach (rank world grad-arena param-arena handles-exchange)
  (with-foreign-objects ((mine :uint8 192) (grads :uint8 (* 64 8)) (params :uint8 (* 64 8)) (sigs :uint8 (* 64 8)))
    (p2p-export (device-vec-ptr (tensor-vec grad-arena)) mine)
    (p2p-export (device-vec-ptr (tensor-vec param-arena)) (inc-pointer mine 64))
    (p2p-signal-export (inc-pointer mine 128))
    (let* ((blob (make-array 192 :element-type '(unsigned-byte 8)))
	   (all  (progn (dotimes (i 192) (setf (aref blob i) (mem-aref mine :uint8 i)))
			(funcall handles-exchange blob))))
      (loop for r from 0 below world
	    for h in all
	    do (dotimes (i 64)
		 (setf (mem-aref grads  :uint8 (+ (* 64 r) i)) (aref h i)
		       (mem-aref params :uint8 (+ (* 64 r) i)) (aref h (+ 64 i))
		       (mem-aref sigs   :uint8 (+ (* 64 r) i)) (aref h (+ 128 i)))))
      (p2p-attach rank world (device-vec-ptr (tensor-vec grad-arena)) (device-vec-ptr (tensor-vec param-arena))
		  grads params sigs))))

(defun cuda-p2p-adam-step (param-arena m-arena v-arena n-floats lr-t-device &key (beta1 0.9) (beta2 0.999) (eps 1e-7))
  "After `backward` on every rank: the whole data-parallel update in one launch."
  (p2p-allreduce-adam-step (device-vec-ptr (tensor-vec param-arena)) 0 0
			   (device-vec-ptr (tensor-vec m-arena)) 0 (device-vec-ptr (tensor-vec v-arena)) 0
			   n-floats 0.0 (device-vec-ptr (tensor-vec lr-t-device)) beta1 beta2 eps))
