;;;; cffi.lisp — CFFI bindings of include/waffe2cuda.h (1:1), loader and status -> condition.
;;;; Counterpart of `cl-waffe2-simd/api.lisp:63-274` + `cl-waffe2-simd/shared-object.lisp:3-12`.
(in-package :cl-waffe2/backends.cuda)

(define-foreign-library libwaffe2cuda
  (t (:default "libwaffe2cuda")))

(defconstant +max-rank+ 8)

(defcstruct wf2cu-view
  (base     :pointer)
  (offset   :int64)
  (rank     :int32)
  (reserved :int32)
  (shape    :int64 :count 8)
  (stride   :int64 :count 8))

(define-condition cuda-backend-error (error)
  ((code :initarg :code) (message :initarg :message))
  (:report (lambda (c s)
	     (format s "libwaffe2cuda: status ~a: ~a" (slot-value c 'code) (slot-value c 'message)))))

(defcfun ("wf2cu_last_error" %last-error) :string)

(defmacro defwf2 (name (&rest args))
  "Bind int wf2cu_<name>(args...) as %<name>, and <name> which signals on a non-zero status.
The VM wraps any error with the instruction context (`source/vm/vm.lisp:230-256`)."
  (let* ((cname (format nil "wf2cu_~a" (substitute #\_ #\- (string-downcase (symbol-name name)))))
	 (raw   (intern (format nil "%~a" name)))
	 (vars  (mapcar #'car args)))
    `(progn
       (defcfun (,cname ,raw) :int ,@args)
       (declaim (inline ,name))
       (defun ,name ,vars
	 (let ((rc (,raw ,@vars)))
	   (unless (zerop rc)
	     (error 'cuda-backend-error :code rc :message (%last-error)))
	   rc)))))

;; runtime
(defwf2 init ((device :int)))
(defwf2 shutdown ())
(defwf2 sync ())
(defwf2 device-name ((buf :pointer) (buflen :size)))
(defwf2 launch-count ((out :pointer)))
(defwf2 device-count ((out :pointer)))
(defwf2 stream ((out :pointer)))
(defwf2 set-stream ((stream :pointer)))
(defwf2 timer-start ())
(defwf2 timer-stop ((out-ms :pointer)))
(defwf2 l2-flush ())
(defwf2 mem-in-use ((out :pointer)))
;; storage
(defwf2 free ((dev :pointer)))
(defwf2 host-alloc ((bytes :size) (out :pointer)))
(defwf2 host-free ((host :pointer)))
(defwf2 malloc ((bytes :size) (out :pointer)))
(defwf2 retain ((dev :pointer)))
(defwf2 release ((dev :pointer)))
(defwf2 memset0 ((dev :pointer) (offset-bytes :size) (bytes :size)))
(defwf2 h2d ((dev :pointer) (offset-bytes :size) (host :pointer) (bytes :size)))
(defwf2 d2h ((dev :pointer) (offset-bytes :size) (host :pointer) (bytes :size)))
(defwf2 d2h-async ((dev :pointer) (offset-bytes :size) (host :pointer) (bytes :size)))
(defwf2 d2d ((dst :pointer) (dst-off :size) (src :pointer) (src-off :size) (bytes :size)))
(defwf2 get-f32 ((dev :pointer) (index :int64) (out :pointer)))
(defwf2 set-f32 ((dev :pointer) (index :int64) (value :float)))
;; graphs
(defwf2 graph-begin ())
(defwf2 graph-end ((out-exec :pointer)))
(defwf2 graph-launch ((exec :pointer)))
(defwf2 graph-destroy ((exec :pointer)))
;; side branches of a compiled step (fork/join edges inside a captured graph)
(defwf2 branch-begin ((b :int)))
(defwf2 branch-end ((b :int)))
(defwf2 branch-join ((b :int)))
;; element-wise
(defwf2 ew-binary ((op :int) (dst :pointer) (src :pointer)))
(defwf2 ew-binary3 ((op :int) (out :pointer) (a :pointer) (b :pointer)))
(defwf2 ew-scalar ((op :int) (dst :pointer) (scalar :float)))
(defwf2 fill ((dst :pointer) (value :float)))
(defwf2 ew-unary ((fn :int) (src :pointer) (dst :pointer)))
(defwf2 pow ((src :pointer) (dst :pointer) (n :float)))
(defwf2 cmp ((op :int) (x :pointer) (y :pointer) (scal :float) (dst :pointer) (then :float) (else :float)))
(defwf2 fused ((program :pointer) (n-instr :int) (views :pointer) (nviews :int)))
(defwf2 fused-set-mode ((mode :int)))
(defwf2 fused-last-path ((out :pointer)))
(defwf2 fused-catalogue-size ((out :pointer)))
;; reductions / composites
(defwf2 reduce-sum ((x :pointer) (out :pointer) (accumulate :int) (divisor :float)))
;; no counterpart in the reference: allocation-free `sum :axis t` (fixed device-array scratch), see include/waffe2cuda.h
(defwf2 sum-all-noalloc ((x :pointer) (x-off :int64) (n :int64) (out :pointer) (out-off :int64) (accumulate :int) (divisor :float)))
(defwf2 softmax ((x :pointer) (out :pointer) (axis :int) (avoid-overflow :int)))
(defwf2 softmax-ex ((x :pointer) (out :pointer) (axis :int) (avoid-overflow :int) (pre-divisor :float)))
(defwf2 reduce-maxmin ((is-min :int) (x :pointer) (out :pointer)))
(defwf2 argmaxmin ((is-min :int) (x :pointer) (out :pointer)))
;; MatmulNode
(defwf2 gemm ((trans-a :int) (trans-b :int) (m :int64) (n :int64) (k :int64)
	      (a :pointer) (a-off :int64) (lda :int64)
	      (b :pointer) (b-off :int64) (ldb :int64)
	      (c :pointer) (c-off :int64) (ldc :int64)
	      (batch :int64) (stride-a :int64) (stride-b :int64) (stride-c :int64)))
;; MatMulNode + the node after it (bias add, !relu, !relu backward mask) in the epilogue
(defwf2 gemm-ex ((trans-a :int) (trans-b :int) (m :int64) (n :int64) (k :int64)
		 (a :pointer) (a-off :int64) (lda :int64)
		 (b :pointer) (b-off :int64) (ldb :int64)
		 (c :pointer) (c-off :int64) (ldc :int64)
		 (bias :pointer) (bias-off :int64) (act :int)
		 (mask :pointer) (mask-off :int64) (ldmask :int64)))
(defwf2 gemm-ex2 ((trans-a :int) (trans-b :int) (m :int64) (n :int64) (k :int64)
		  (a :pointer) (a-off :int64) (lda :int64) (b :pointer) (b-off :int64) (ldb :int64)
		  (c :pointer) (c-off :int64) (ldc :int64) (bias :pointer) (bias-off :int64) (act :int)
		  (residual :pointer) (res-off :int64) (ldres :int64)))
;; (!sum (softmax-cross-entropy x labels)) forward / sce-diff-static backward
(defwf2 sce-sum-forward ((x :pointer) (x-off :int64) (labels :pointer) (l-off :int64) (z :pointer) (z-off :int64)
			 (loss :pointer) (loss-off :int64) (rows :int64) (cols :int64) (delta :float)))
(defwf2 sce-backward ((z :pointer) (z-off :int64) (labels :pointer) (l-off :int64) (dy :pointer) (dy-off :int64)
		      (dx :pointer) (dx-off :int64) (rows :int64) (cols :int64) (accumulate :int)))
(defwf2 gemm-set-mode ((mode :int)))
(defwf2 gemm-set-correction ((kind :int)))
(defwf2 emit-operand-planes ((on :int)))
(defwf2 gemm-last-path ((out :pointer)))
(defwf2 gemm-const-operand ((base :pointer) (offset :int64) (count :int64) (constant :int)))
(defwf2 gemm-const-stats ((entries :pointer) (bytes :pointer)))
;; optimizers
(defwf2 sgd-step ((param :pointer) (p-off :int64) (grad :pointer) (g-off :int64) (n :int64) (lr :float)))
(defwf2 adam-step ((param :pointer) (p-off :int64) (grad :pointer) (g-off :int64)
		   (m :pointer) (m-off :int64) (v :pointer) (v-off :int64) (n :int64)
		   (lr-t :float) (lr-t-dev :pointer) (beta1 :float) (beta2 :float) (eps :float) (grad-scale :float)))
(defwf2 embedding-gather ((ctx :pointer) (batch :int64) (seq :int64) (wte :pointer) (wpe :pointer) (dim :int64) (out :pointer)))
;; the GPT-2 example's SelfAttention as two launches (examples/gpt-2/utils.lisp:204-244)
(defwf2 attn-heads-softmax ((qkv :pointer) (qkv-off :int64) (batch :int64) (seq :int64) (n-head :int64) (head-dim :int64)
			    (divisor :float) (p :pointer) (p-off :int64)))
(defwf2 attn-pv ((p :pointer) (p-off :int64) (qkv :pointer) (qkv-off :int64) (batch :int64) (seq :int64) (n-head :int64)
		 (head-dim :int64) (out :pointer) (out-off :int64)))
(defwf2 attn-fused ((qkv :pointer) (qkv-off :int64) (batch :int64) (seq :int64) (n-head :int64) (head-dim :int64)
		    (divisor :float) (out :pointer) (out-off :int64)))
(defwf2 layernorm ((x :pointer) (x-off :int64) (g :pointer) (g-off :int64) (b :pointer) (b-off :int64)
		   (out :pointer) (out-off :int64) (rows :int64) (cols :int64) (eps :float)))
;; unfold / Conv2D / pooling (csrc/conv.cu)
(defcstruct wf2cu-unfold-geom
  (n :int64) (c :int64) (h :int64) (w :int64)
  (h-out :int64) (w-out :int64) (k-h :int64) (k-w :int64)
  (pad-h :int64) (pad-w :int64) (stride-h :int64) (stride-w :int64) (dilation-h :int64) (dilation-w :int64))
(defwf2 im2col ((col :pointer) (col-off :int64) (im :pointer) (im-off :int64) (geom :pointer)))
(defwf2 col2im ((col :pointer) (col-off :int64) (im :pointer) (im-off :int64) (geom :pointer)))
(defwf2 unfold ((x :pointer) (rows :pointer) (rows-off :int64) (geom :pointer)))
(defwf2 fold ((drows :pointer) (drows-off :int64) (dx :pointer) (geom :pointer) (accumulate :int)))
(defwf2 pool2d ((mode :int) (x :pointer) (out :pointer) (geom :pointer)))
(defwf2 pool2d-backward ((mode :int) (x :pointer) (out :pointer) (dout :pointer) (dx :pointer) (geom :pointer)
			 (accumulate :int) (relu-input :int)))
;; multi-GPU
(defwf2 comm-unique-id ((out128 :pointer)))
(defwf2 comm-init ((rank :int) (world :int) (id128 :pointer)))
(defwf2 allreduce-sum-f32 ((buf :pointer) (offset :int64) (count :int64)))
(defwf2 comm-destroy ())
;; exchange + Adam as one kernel over NVLink peer memory (CUDA IPC, one process per GPU)
(defwf2 p2p-export ((dev-alloc :pointer) (handle64 :pointer)))
(defwf2 p2p-signal-export ((handle64 :pointer)))
(defwf2 p2p-attach ((rank :int) (world :int) (local-grad :pointer) (local-param :pointer)
		    (grad-handles :pointer) (param-handles :pointer) (signal-handles :pointer)))
(defwf2 p2p-allreduce-adam-step ((param :pointer) (p-off :int64) (g-off :int64) (m :pointer) (m-off :int64)
				 (v :pointer) (v-off :int64) (n :int64) (lr-t :float) (lr-t-dev :pointer)
				 (beta1 :float) (beta2 :float) (eps :float)))
(defwf2 p2p-status ((timed-out :pointer)))
(defwf2 p2p-detach ())

;; enums of waffe2cuda.h
(defconstant +pool-max+ 0) (defconstant +pool-avg+ 1)
(defconstant +add+ 0) (defconstant +sub+ 1) (defconstant +mul+ 2) (defconstant +div+ 3) (defconstant +copy+ 4)
(defparameter *unary-codes*
  '(:sin 0 :cos 1 :tan 2 :asin 3 :acos 4 :atan 5 :sinh 6 :cosh 7 :tanh 8 :abs 9 :exp 10 :loge 11
    :log2 12 :log10 13 :log1p 14 :sqrt 15 :cbrt 16 :sign 17 :square 18 :neg 19 :recip 20))
(defun cmp-code (condition)
  "Same mapping as `cond->fname` (`source/backends/cpu/logical.lisp:28-35`)."
  (cond ((eql condition #'=) 0) ((eql condition #'<) 1) ((eql condition #'<=) 2)
	((eql condition #'>) 3) ((eql condition #'>=) 4)
	(t (error "CUDATensor: unsupported comparison ~a" condition))))

(defun load-cuda-backend (&key (device 0))
  "Load the library and select the device.  On failure: warn and leave the backend disabled, as
`cl-waffe2-simd/shared-object.lisp:3-12` does."
  (handler-case
      (progn
	(use-foreign-library libwaffe2cuda)
	(init device)
	(setf *cuda-available-p* t))
    (error (c)
      (warn "cl-waffe2/backends.cuda: CUDATensor is disabled (~a)" c)
      (setf *cuda-available-p* nil))))
