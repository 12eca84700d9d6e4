ssh test@127.0.0.1 "kill -9 9013"
ssh test@127.0.0.1 "kill -9 9014"
ssh test@127.0.0.1 "kill -9 9015"
ssh test@127.0.0.1 "kill -9 9016"
kill -9 8873
