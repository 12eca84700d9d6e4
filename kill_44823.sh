ssh test@127.0.0.1 "kill -9 9509"
kill -9 8873
