ssh test@127.0.0.1 "kill -9 9203"
ssh test@127.0.0.1 "kill -9 9204"
ssh test@127.0.0.1 "kill -9 9205"
kill -9 8873
